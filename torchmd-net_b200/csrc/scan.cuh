// scan.cuh -- device-wide exclusive scan over int32 (see scan.cu).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>
size_t tmdn_scan_tmp_ints(int64_t n_out);
int tmdn_scan_exclusive_i32(const int* in, int* out, int64_t n_in, int* tmp, cudaStream_t stream);
