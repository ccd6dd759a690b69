// General-shape MLP kernels, hidden = (64, 64): see mlp_general.cuh / mlp_general_inst.cuh
#define GEN_H 64
#include "mlp_general_inst.cuh"
