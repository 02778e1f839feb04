#include "kernels.cuh"
FT_DEFINE_GROUP_LAUNCH(2)
