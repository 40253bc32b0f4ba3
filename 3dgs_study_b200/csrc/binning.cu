// binning.cu -- tile binning: inclusive scan of tiles_touched, duplicateWithKeys (64-bit
// tile|depth keys emitted in Gaussian-index order, row-major over each tile rectangle), stable
// radix sort of the (key, Gaussian) pairs and identifyTileRanges.  HBM-bound integer work.
//
// Stands in for the binning stages of _C.rasterize_gaussians named by BASELINE.json's north_star
// ("duplicateWithKeys pass emitting 64-bit tile|depth keys, then a radix sort and identifyTileRanges").
#include <cub/cub.cuh>

#include "splat_internal.cuh"

namespace b200splat {

size_t scan_temp_bytes(int P) {
    size_t bytes = 0;
    cub::DeviceScan::InclusiveSum(nullptr, bytes, (uint32_t*)nullptr, (uint32_t*)nullptr, P > 0 ? P : 1);
    return bytes;
}

size_t sort_temp_bytes(int R, int end_bit) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (uint64_t*)nullptr, (uint64_t*)nullptr, (uint32_t*)nullptr,
                                    (uint32_t*)nullptr, R > 0 ? R : 1, 0, end_bit);
    return bytes;
}

// number of key bits that can differ: 32 depth bits + bits of the largest tile id
int sort_end_bit(int W, int H) {
    unsigned tiles = (unsigned)(((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE));
    int bits = 0;
    while ((tiles >> bits) != 0 && bits < 32) bits++;
    return 32 + bits;
}

GeomView carve_geom(void* base, int P) {
    GeomView g;
    size_t off = 0;
    char* b = (char*)base;
    size_t n = (size_t)(P > 0 ? P : 1);
    g.depths = (float*)(b + off); off += align256(n * sizeof(float));
    g.rec = (float4*)(b + off); off += align256(n * 3 * sizeof(float4));
    g.tiles = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    g.offsets = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    g.scan_tmp_bytes = scan_temp_bytes(P);
    g.scan_tmp = (void*)(b + off); off += align256(g.scan_tmp_bytes);
    g.total = off;
    return g;
}

BinView carve_binning(void* base, int R, int end_bit) {
    BinView v;
    size_t off = 0;
    char* b = (char*)base;
    size_t n = (size_t)(R > 0 ? R : 1);
    v.keys_unsorted = (uint64_t*)(b + off); off += align256(n * sizeof(uint64_t));
    v.keys_sorted = (uint64_t*)(b + off); off += align256(n * sizeof(uint64_t));
    v.vals_unsorted = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    v.vals_sorted = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    v.sort_tmp_bytes = sort_temp_bytes(R, end_bit);
    v.sort_tmp = (void*)(b + off); off += align256(v.sort_tmp_bytes);
    v.total = off;
    return v;
}

ImgView carve_image(void* base, int W, int H) {
    ImgView v;
    size_t off = 0;
    char* b = (char*)base;
    size_t T = (size_t)((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
    size_t N = (size_t)W * H;
    v.ranges = (int2*)(b + off); off += align256((T > 0 ? T : 1) * sizeof(int2));
    v.final_T = (float*)(b + off); off += align256((N > 0 ? N : 1) * sizeof(float));
    v.n_contrib = (uint32_t*)(b + off); off += align256((N > 0 ? N : 1) * sizeof(uint32_t));
    v.counters = (unsigned int*)(b + off); off += align256(64 * sizeof(unsigned int));
    v.total = off;
    return v;
}

namespace {

// One thread per Gaussian; keys are written at [offsets[i-1], offsets[i]) in row-major tile order so
// that a stable sort leaves equal (tile, depth) keys in Gaussian-index order.
__global__ void __launch_bounds__(256) duplicate_with_keys_kernel(int P, int gx, int gy, const int32_t* __restrict__ radii,
                                                                  const float4* __restrict__ rec,
                                                                  const float* __restrict__ depths,
                                                                  const uint32_t* __restrict__ offsets,
                                                                  uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int r = radii[i];
    if (r <= 0) return;
    uint32_t off = (i == 0) ? 0u : offsets[i - 1];
    const float4 a = __ldg(rec + 3 * (size_t)i);
    int minx, miny, maxx, maxy;
    get_rect(a.x, a.y, r, gx, gy, minx, miny, maxx, maxy);
    const uint64_t dbits = (uint64_t)__float_as_uint(depths[i]);
    for (int y = miny; y < maxy; y++) {
        for (int x = minx; x < maxx; x++) {
            keys[off] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | dbits;
            vals[off] = (uint32_t)i;
            off++;
        }
    }
}

__global__ void __launch_bounds__(256) identify_tile_ranges_kernel(int R, const uint64_t* __restrict__ keys,
                                                                   int2* __restrict__ ranges) {
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

}  // namespace

cudaError_t launch_scan(GeomView gv, int P, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    size_t bytes = gv.scan_tmp_bytes;
    return cub::DeviceScan::InclusiveSum(gv.scan_tmp, bytes, gv.tiles, gv.offsets, P, s);
}

cudaError_t launch_duplicate(GeomView gv, const int32_t* radii, int P, int W, int H, BinView bv, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    duplicate_with_keys_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, gx, gy, radii, gv.rec, gv.depths, gv.offsets,
                                                             bv.keys_unsorted, bv.vals_unsorted);
    return cudaGetLastError();
}

cudaError_t launch_sort(BinView bv, int R, int end_bit, cudaStream_t s) {
    if (R == 0) return cudaSuccess;
    size_t bytes = bv.sort_tmp_bytes;
    return cub::DeviceRadixSort::SortPairs(bv.sort_tmp, bytes, bv.keys_unsorted, bv.keys_sorted, bv.vals_unsorted,
                                           bv.vals_sorted, R, 0, end_bit, s);
}

cudaError_t launch_ranges(BinView bv, int R, ImgView iv, int W, int H, cudaStream_t s) {
    const size_t T = (size_t)((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
    cudaError_t e = cudaMemsetAsync(iv.ranges, 0, T * sizeof(int2), s);
    if (e != cudaSuccess || R == 0) return e;
    identify_tile_ranges_kernel<<<(R + 255) / 256, 256, 0, s>>>(R, bv.keys_sorted, iv.ranges);
    return cudaGetLastError();
}

}  // namespace b200splat
