// cx_launch.cuh — launch bookkeeping shared by the translation units of libcotix_b200.so
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

// defined in abi.cu
extern "C" int cx_sm_count(void);
extern "C" void cx_count_launch(void);
extern "C" int cx_check_launch(const char* what);
extern "C" void cx_set_error(const char* fmt, ...);
extern "C" void cx_count_launch_n(int64_t n);
