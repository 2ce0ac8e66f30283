// Scan encoder (state2costmap + FasterNet forward), see encoder.cu.
#pragma once
#include "gemm.cuh"
