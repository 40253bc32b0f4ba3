// binning_classic.cu -- COMPARATOR ONLY, never part of libb200splat.so.
//
// The reference's own CUDA rasterizer is absent from the snapshot (/root/reference/.SUBMODULES.json:16-22), so
// the ">= 2x the reference rasterizer" target cannot be measured against it.  This file and render_classic.cu
// build a clearly-labelled BASELINE-ALGORITHM VARIANT of this library (tools/build_classic.sh ->
// variants/libb200splat_classic.so, selected with B200SPLAT_LIB / `bench.py --impl classic`): the textbook
// formulation BASELINE.json's north_star describes as the thing being replaced -- scan of tiles touched in
// Gaussian order with a BLOCKING read-back of the pair count, one-thread-per-Gaussian duplicateWithKeys
// emitting 64-bit tile|depth keys, one radix sort over all 32 + log2(T) key bits, one-thread-per-pair
// identifyTileRanges, one 16x16 block per tile with one pixel per thread, and a backward pass that issues nine
// global float atomics per (pixel, Gaussian).  It shares api.cu and the per-Gaussian kernels with the product
// (which flatters it) and is written from the published algorithm description, not from any reference source.
// It is a stand-in, never presented as the reference.
#include <cub/cub.cuh>

#include "../splat_internal.cuh"

namespace b200splat {

namespace {

struct TilesInIndexOrder {
    const uint2* rect;
    __device__ __forceinline__ uint32_t operator()(uint32_t i) const {
        const uint32_t wh = rect[i].y;
        return (wh & 0xffffu) * (wh >> 16);
    }
};
using TilesIt = cub::TransformInputIterator<uint32_t, TilesInIndexOrder, cub::CountingInputIterator<uint32_t>>;

size_t scan_temp_bytes(int P) {
    size_t a = 0;
    TilesIt it(cub::CountingInputIterator<uint32_t>(0), TilesInIndexOrder{nullptr});
    cub::DeviceScan::InclusiveSum(nullptr, a, it, (uint32_t*)nullptr, P > 0 ? P : 1);
    return a;
}
int key_end_bit(int num_tiles) {
    int bits = 1;
    while (bits < 32 && ((unsigned)(num_tiles - 1) >> bits) != 0) bits++;
    return 32 + bits;
}
size_t sort_temp_bytes(int R, int num_tiles) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (uint64_t*)nullptr, (uint64_t*)nullptr, (uint32_t*)nullptr,
                                    (uint32_t*)nullptr, R > 0 ? R : 1, 0, key_end_bit(num_tiles));
    return bytes;
}

}  // namespace

GeomView carve_geom(void* base, int P) {
    GeomView g;
    size_t off = 0;
    char* b = (char*)base;
    size_t n = (size_t)(P > 0 ? P : 1);
    g.depths = (float*)(b + off); off += align256(n * sizeof(float));
    g.rec = (float4*)(b + off); off += align256(n * 3 * sizeof(float4));
    g.rect = (uint2*)(b + off); off += align256(n * sizeof(uint2));
    g.iota = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    g.depth_sorted = nullptr; g.order = nullptr;
    g.offsets = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    g.scan_tmp_bytes = scan_temp_bytes(P);
    g.scan_tmp = (void*)(b + off); off += align256(g.scan_tmp_bytes);
    g.dsort_tmp = nullptr; g.dsort_tmp_bytes = 0;
    g.total = off;
    return g;
}

BinView carve_binning(void* base, int R, int num_tiles) {
    BinView v;
    size_t off = 0;
    char* b = (char*)base;
    size_t n = (size_t)(R > 0 ? R : 1);
    v.key_bytes = 8;
    v.keys_unsorted = (void*)(b + off); off += align256(n * 8);
    v.keys_sorted = (void*)(b + off); off += align256(n * 8);
    v.vals_unsorted = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    v.vals_sorted = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    v.sort_tmp_bytes = sort_temp_bytes(R, num_tiles);
    v.sort_tmp = (void*)(b + off); off += align256(v.sort_tmp_bytes);
    v.total = off;
    return v;
}

ImgView carve_image(void* base, int W, int H) {
    ImgView v;
    size_t off = 0;
    char* b = (char*)base;
    size_t T = (size_t)tiles_of(W, H);
    size_t N = (size_t)W * H;
    v.ranges = (int2*)(b + off); off += align256((T > 0 ? T : 1) * sizeof(int2));
    v.final_T = (float*)(b + off); off += align256((N > 0 ? N : 1) * sizeof(float));
    v.n_contrib = (uint32_t*)(b + off); off += align256((N > 0 ? N : 1) * sizeof(uint32_t));
    v.counters = (unsigned int*)(b + off); off += align256(64 * sizeof(unsigned int));
    v.tile_order = nullptr;
    v.total = off;
    return v;
}

namespace {

__global__ void __launch_bounds__(256) duplicate_with_keys_kernel(int P, int gx, const uint2* __restrict__ rect,
                                                                  const float* __restrict__ depths,
                                                                  const uint32_t* __restrict__ offsets,
                                                                  uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint2 r = rect[i];
    const int w = r.y & 0xffffu, h = r.y >> 16;
    if (w * h == 0) return;
    uint32_t off = (i == 0) ? 0u : offsets[i - 1];
    const int minx = r.x & 0xffffu, miny = r.x >> 16;
    const uint64_t dbits = (uint64_t)__float_as_uint(depths[i]);
    for (int y = miny; y < miny + h; y++)
        for (int x = minx; x < minx + w; x++) {
            keys[off] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | dbits;
            vals[off] = (uint32_t)i;
            off++;
        }
}

__global__ void __launch_bounds__(256) identify_tile_ranges_kernel(int R, const uint64_t* __restrict__ keys, int2* __restrict__ ranges) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= R) return;
    const uint32_t t = (uint32_t)(keys[j] >> 32);
    if (j == 0) ranges[t].x = 0;
    else {
        const uint32_t tp = (uint32_t)(keys[j - 1] >> 32);
        if (tp != t) { ranges[tp].y = j; ranges[t].x = j; }
    }
    if (j == R - 1) ranges[t].y = R;
}

__global__ void export_tiles_kernel(int P, const uint2* __restrict__ rect, uint32_t* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint32_t wh = rect[i].y;
    out[i] = (wh & 0xffffu) * (wh >> 16);
}

}  // namespace

// scan in Gaussian order, then a BLOCKING synchronisation (the classic formulation reads R back here)
cudaError_t launch_depth_order(GeomView gv, int P, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    TilesIt it(cub::CountingInputIterator<uint32_t>(0), TilesInIndexOrder{gv.rect});
    size_t bytes = gv.scan_tmp_bytes;
    cudaError_t e = cub::DeviceScan::InclusiveSum(gv.scan_tmp, bytes, it, gv.offsets, P, s);
    if (e != cudaSuccess) return e;
    return cudaStreamSynchronize(s);
}

cudaError_t launch_emit(GeomView gv, int P, int W, int H, BinView bv, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    const int gx = (W + TILE - 1) / TILE;
    duplicate_with_keys_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, gx, gv.rect, gv.depths, gv.offsets, (uint64_t*)bv.keys_unsorted,
                                                             bv.vals_unsorted);
    return cudaGetLastError();
}

cudaError_t launch_tile_sort(BinView bv, int R, int num_tiles, cudaStream_t s) {
    if (R == 0) return cudaSuccess;
    size_t bytes = bv.sort_tmp_bytes;
    return cub::DeviceRadixSort::SortPairs(bv.sort_tmp, bytes, (const uint64_t*)bv.keys_unsorted, (uint64_t*)bv.keys_sorted,
                                           bv.vals_unsorted, bv.vals_sorted, R, 0, key_end_bit(num_tiles), s);
}

cudaError_t launch_ranges(BinView bv, int R, ImgView iv, int W, int H, cudaStream_t s) {
    const size_t T = (size_t)tiles_of(W, H);
    if (T == 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(iv.ranges, 0, T * sizeof(int2), s);
    if (e != cudaSuccess || R == 0) return e;
    identify_tile_ranges_kernel<<<(R + 255) / 256, 256, 0, s>>>(R, (const uint64_t*)bv.keys_sorted, iv.ranges);
    return cudaGetLastError();
}

cudaError_t launch_export_binning(int P, int R, GeomView gv, BinView bv, const b200splat_state_out& o, cudaStream_t s) {
    if (o.keys_sorted && R > 0) {
        cudaError_t e = cudaMemcpyAsync(o.keys_sorted, bv.keys_sorted, sizeof(uint64_t) * (size_t)R, cudaMemcpyDeviceToDevice, s);
        if (e != cudaSuccess) return e;
    }
    if (o.tiles_touched && P > 0) export_tiles_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, gv.rect, o.tiles_touched);
    if (o.point_offsets && P > 0) {
        cudaError_t e = cudaMemcpyAsync(o.point_offsets, gv.offsets, sizeof(uint32_t) * (size_t)P, cudaMemcpyDeviceToDevice, s);
        if (e != cudaSuccess) return e;
    }
    return cudaGetLastError();
}

}  // namespace b200splat
