// General-shape persistent training loop, hidden = (32, 32): see train_loop_general_inst.cuh
#define GEN_H 32
#include "train_loop_general_inst.cuh"
