// cg_store.cuh -- maintenance kernels of the device-resident scrape history.
//
// The reference keeps no history: every attempt asks Prometheus for "now" (mlflow_operator.py
// :302-303, then sleeps 10 s or 60 s, :340, :352) and Prometheus' TSDB holds the hour.  Here the
// hour lives in HBM in the library's SoA layout, so an evaluation costs one new scrape over PCIe,
// not the whole history: rows have some slack after the newest sample; cg_store_append writes the
// next scrape into column `column` of every row, and when the slack is used up cg_store_shift moves
// every row left so that the oldest samples fall off.
#pragma once
#include <stdint.h>
#include "../../include/canary_gate.h"

namespace cg {

// one thread per series: rows[series][column] = fresh[series]
template <typename T>
__global__ void store_append_kernel(T* __restrict__ rows, const T* __restrict__ fresh, int64_t n_series,
                                    int row_pitch, int column) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n_series; i += (int64_t)gridDim.x * blockDim.x)
        rows[i * row_pitch + column] = fresh[i];
}

// one CTA per row at a time: rows[r][0 .. count) = rows[r][shift .. shift + count), staged through
// shared memory so that overlapping source and destination are safe
template <typename T>
__global__ void store_shift_kernel(T* __restrict__ rows, int64_t n_rows, int row_pitch, int shift, int count) {
    extern __shared__ __align__(16) unsigned char cg_store_smem[];
    T* stage = reinterpret_cast<T*>(cg_store_smem);
    for (int64_t r = blockIdx.x; r < n_rows; r += gridDim.x) {
        T* row = rows + r * row_pitch;
        for (int i = threadIdx.x; i < count; i += blockDim.x) stage[i] = row[shift + i];
        __syncthreads();
        for (int i = threadIdx.x; i < count; i += blockDim.x) row[i] = stage[i];
        __syncthreads();
    }
}

}  // namespace cg
