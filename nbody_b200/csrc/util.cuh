/*
 * util.cuh -- failure convention of the CUDA side.
 *
 * Mirrors the reference's ASSERT / ASSERT_VK (ref src/lib/util.h:17-29,47-60): there are no
 * return codes; a failed check prints "file:line [func] ..." to stderr and abort()s.
 */
#pragma once

#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#define NB_FAIL(...)                                                        \
    do {                                                                    \
        fprintf(stderr, "%s:%d [%s] ", __FILE__, __LINE__, __func__);       \
        fprintf(stderr, __VA_ARGS__);                                       \
        fprintf(stderr, "\n");                                              \
        abort();                                                            \
    } while (0)

#define NB_ASSERT(COND, ...)              \
    do {                                  \
        if (!(COND)) NB_FAIL(__VA_ARGS__); \
    } while (0)

#define NB_CUDA(X)                                                                         \
    do {                                                                                   \
        cudaError_t nb_cuda_err_ = (X);                                                    \
        if (nb_cuda_err_ != cudaSuccess)                                                   \
            NB_FAIL("cudaError = %d, str = %s: %s", (int)nb_cuda_err_,                      \
                    cudaGetErrorString(nb_cuda_err_), #X);                                 \
    } while (0)
