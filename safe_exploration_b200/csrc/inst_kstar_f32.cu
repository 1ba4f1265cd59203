#define SE_T float
#include "inst_kstar.cuh"
