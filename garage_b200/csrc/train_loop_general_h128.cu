// General-shape persistent training loop, hidden = (128, 128): see train_loop_general_inst.cuh
#define GEN_H 128
#include "train_loop_general_inst.cuh"
