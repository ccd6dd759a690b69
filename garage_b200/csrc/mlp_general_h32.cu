// General-shape MLP kernels, hidden = (32, 32): see mlp_general.cuh / mlp_general_inst.cuh
#define GEN_H 32
#include "mlp_general_inst.cuh"
