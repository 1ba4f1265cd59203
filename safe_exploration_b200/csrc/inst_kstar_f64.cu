#define SE_T double
#include "inst_kstar.cuh"
