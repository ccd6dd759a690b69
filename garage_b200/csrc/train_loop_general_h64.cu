// General-shape persistent training loop, hidden = (64, 64): see train_loop_general_inst.cuh
#define GEN_H 64
#include "train_loop_general_inst.cuh"
