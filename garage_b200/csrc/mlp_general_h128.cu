// General-shape MLP kernels, hidden = (128, 128): see mlp_general.cuh / mlp_general_inst.cuh
#define GEN_H 128
#include "mlp_general_inst.cuh"
