// binning.cu -- tile binning, sm_100a.  HBM-bound integer work.
//
// Stands in for the binning stages of _C.rasterize_gaussians named by BASELINE.json's north_star
// ("duplicateWithKeys pass emitting 64-bit tile|depth keys, then a radix sort and identifyTileRanges").
// The RESULT is bit-identical to that formulation -- every tile's list is ordered by (depth bits,
// Gaussian id), exactly what a stable sort of (tile << 32 | depth) keys emitted in Gaussian order
// yields -- but the traffic is a fraction of it (see splat_internal.cuh "Binning order"):
//
//   1. depth order   stable radix sort of the P (depth bits, id) pairs      4 passes x 16 B x P
//   2. offsets       inclusive scan of tiles touched, in depth order          gather + 4 B x P
//   3. emit          block-cooperative duplicateWithKeys: a block owns 256 consecutive Gaussians of the
//                    depth order and therefore a CONTIGUOUS output range; consecutive threads write
//                    consecutive (tile key, id) slots (binary search in the block's offsets), so the
//                    writes are fully coalesced and a huge splat cannot serialise one thread
//   4. tile sort     stable radix sort on the tile bits only                  2 passes x 12 B x R
//   5. ranges        identifyTileRanges on the sorted tile keys
//
// The 64-bit keys are never materialised on the hot path; b200splat_export_state rebuilds them for
// the parity tests from (sorted tile key, depth of the sorted id).
#include <cub/cub.cuh>

#include "splat_internal.cuh"

namespace b200splat {

namespace {

// tiles touched by the j-th Gaussian of the depth order
struct TilesInDepthOrder {
    const uint32_t* order;
    const uint2* rect;
    __device__ __forceinline__ uint32_t operator()(uint32_t j) const {
        const uint32_t wh = __ldg(&rect[__ldg(order + j)].y);
        return (wh & 0xffffu) * (wh >> 16);
    }
};
// tiles touched by Gaussian i (index order; export only)
struct TilesInIndexOrder {
    const uint2* rect;
    __device__ __forceinline__ uint32_t operator()(uint32_t i) const {
        const uint32_t wh = rect[i].y;
        return (wh & 0xffffu) * (wh >> 16);
    }
};
template <typename Op>
using TilesIt = cub::TransformInputIterator<uint32_t, Op, cub::CountingInputIterator<uint32_t>>;

size_t scan_temp_bytes(int P) {
    size_t a = 0, b = 0;
    TilesIt<TilesInDepthOrder> it(cub::CountingInputIterator<uint32_t>(0), TilesInDepthOrder{nullptr, nullptr});
    cub::DeviceScan::InclusiveSum(nullptr, a, it, (uint32_t*)nullptr, P > 0 ? P : 1);
    TilesIt<TilesInIndexOrder> it2(cub::CountingInputIterator<uint32_t>(0), TilesInIndexOrder{nullptr});
    cub::DeviceScan::InclusiveSum(nullptr, b, it2, (uint32_t*)nullptr, P > 0 ? P : 1);
    return a > b ? a : b;
}

size_t depth_sort_temp_bytes(int P) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (uint32_t*)nullptr, (uint32_t*)nullptr, (uint32_t*)nullptr,
                                    (uint32_t*)nullptr, P > 0 ? P : 1, 0, 32);
    return bytes;
}

int tile_bits(int num_tiles) {
    int bits = 1;
    while (bits < 32 && ((unsigned)(num_tiles - 1) >> bits) != 0) bits++;
    return bits;
}

template <typename KeyT>
size_t tile_sort_temp_bytes(int R, int num_tiles) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (KeyT*)nullptr, (KeyT*)nullptr, (uint32_t*)nullptr,
                                    (uint32_t*)nullptr, R > 0 ? R : 1, 0, tile_bits(num_tiles));
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
    g.depth_sorted = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    g.order = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    g.offsets = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    g.scan_tmp_bytes = scan_temp_bytes(P);
    g.scan_tmp = (void*)(b + off); off += align256(g.scan_tmp_bytes);
    g.dsort_tmp_bytes = depth_sort_temp_bytes(P);
    g.dsort_tmp = (void*)(b + off); off += align256(g.dsort_tmp_bytes);
    g.total = off;
    return g;
}

BinView carve_binning(void* base, int R, int num_tiles) {
    BinView v;
    size_t off = 0;
    char* b = (char*)base;
    size_t n = (size_t)(R > 0 ? R : 1);
    v.key_bytes = num_tiles <= 65536 ? 2 : 4;
    v.keys_unsorted = (void*)(b + off); off += align256(n * v.key_bytes);
    v.keys_sorted = (void*)(b + off); off += align256(n * v.key_bytes);
    v.vals_unsorted = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    v.vals_sorted = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    v.sort_tmp_bytes = v.key_bytes == 2 ? tile_sort_temp_bytes<uint16_t>(R, num_tiles) : tile_sort_temp_bytes<uint32_t>(R, num_tiles);
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
    const size_t Tn = T > 0 ? T : 1;
    v.tile_order = (uint32_t*)(b + off); off += align256(Tn * sizeof(uint32_t));
    v.total = off;
    return v;
}

namespace {

constexpr int EMIT_THREADS = 256;

// Block-cooperative duplicateWithKeys.  Block b owns depth-order positions [256 b, 256 b + 256) and the
// output slots [offsets[256 b - 1], offsets[256 b + 255]).  Slot k of Gaussian g (rect minx, miny, w, h)
// is tile (miny + k / w) * gx + minx + k % w: row-major over the rectangle, as the oracle emits.
template <typename KeyT>
__global__ void __launch_bounds__(EMIT_THREADS) emit_keys_kernel(int P, int gx, const uint32_t* __restrict__ order,
                                                                 const uint2* __restrict__ rect,
                                                                 const uint32_t* __restrict__ offsets,
                                                                 KeyT* __restrict__ keys, uint32_t* __restrict__ vals) {
    __shared__ uint32_t s_end[EMIT_THREADS];     // inclusive offsets of this block's Gaussians
    __shared__ uint32_t s_id[EMIT_THREADS];
    __shared__ uint2 s_rect[EMIT_THREADS];
    const int tid = threadIdx.x;
    const int j0 = blockIdx.x * EMIT_THREADS;
    const int j = min(j0 + tid, P - 1);
    const uint32_t id = __ldg(order + j);
    s_id[tid] = id;
    s_rect[tid] = __ldg(rect + id);
    s_end[tid] = __ldg(offsets + j);
    const uint32_t base = j0 == 0 ? 0u : __ldg(offsets + j0 - 1);
    __syncthreads();
    // Each thread owns four consecutive output slots (aligned to 4 so full groups go out as one vector store
    // per array): one binary search and one division for the first slot, then a linear walk.
    const uint32_t end = s_end[EMIT_THREADS - 1];
    for (uint32_t gs = (base & ~3u) + 4u * tid; gs < end; gs += 4u * EMIT_THREADS) {
        const uint32_t s0 = max(gs, base);
        int lo = 0, hi = EMIT_THREADS - 1;          // first g with s_end[g] > s0
#pragma unroll
        for (int it = 0; it < 8; it++) {
            const int mid = (lo + hi) >> 1;
            if (s_end[mid] > s0) hi = mid; else lo = mid + 1;
        }
        int g = lo;
        uint2 r = s_rect[g];
        uint32_t w = r.y & 0xffffu;
        uint32_t gend = s_end[g];
        uint32_t id = s_id[g];
        const uint32_t local = s0 - (gend - w * (r.y >> 16));
        const uint32_t row = local / w;
        uint32_t col = local - row * w;
        uint32_t tile = ((r.x >> 16) + row) * (uint32_t)gx + (r.x & 0xffffu) + col;
        KeyT kk[4];
        uint32_t vv[4];
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const uint32_t sl = gs + e;
            kk[e] = 0; vv[e] = 0;
            if (sl >= s0 && sl < end) {
                while (sl >= gend) {                 // next Gaussian with tiles: restart at its first tile
                    g++;
                    r = s_rect[g]; w = r.y & 0xffffu; gend = s_end[g]; id = s_id[g];
                    col = 0;
                    tile = (r.x >> 16) * (uint32_t)gx + (r.x & 0xffffu);
                }
                kk[e] = (KeyT)tile; vv[e] = id;
                col++; tile++;
                if (col == w) { col = 0; tile += (uint32_t)gx - w; }
            }
        }
        if (gs >= base && gs + 4 <= end) {
            if (sizeof(KeyT) == 2)
                *reinterpret_cast<uint2*>(keys + gs) = make_uint2((uint32_t)kk[0] | ((uint32_t)kk[1] << 16), (uint32_t)kk[2] | ((uint32_t)kk[3] << 16));
            else
                *reinterpret_cast<uint4*>(keys + gs) = make_uint4(kk[0], kk[1], kk[2], kk[3]);
            *reinterpret_cast<uint4*>(vals + gs) = make_uint4(vv[0], vv[1], vv[2], vv[3]);
        } else {
#pragma unroll
            for (int e = 0; e < 4; e++)
                if (gs + e >= s0 && gs + e < end) { keys[gs + e] = kk[e]; vals[gs + e] = vv[e]; }
        }
    }
}

// identifyTileRanges: each thread owns EPT consecutive sorted keys (one 16-byte load) plus the key before them.
template <typename KeyT>
__global__ void __launch_bounds__(256) identify_tile_ranges_kernel(int R, const KeyT* __restrict__ keys,
                                                                   int2* __restrict__ ranges) {
    constexpr int EPT = 16 / sizeof(KeyT);
    const int j0 = (blockIdx.x * blockDim.x + threadIdx.x) * EPT;
    if (j0 >= R) return;
    KeyT k[EPT];
    if (j0 + EPT <= R) {
        *reinterpret_cast<uint4*>(k) = *reinterpret_cast<const uint4*>(keys + j0);
    } else {
#pragma unroll
        for (int e = 0; e < EPT; e++) k[e] = j0 + e < R ? keys[j0 + e] : (KeyT)0;
    }
    uint32_t prev = j0 > 0 ? (uint32_t)keys[j0 - 1] : 0u;
    if (j0 == 0) ranges[(uint32_t)k[0]].x = 0;
#pragma unroll
    for (int e = 0; e < EPT; e++) {
        const int j = j0 + e;
        if (j < R) {
            const uint32_t t = (uint32_t)k[e];
            if (j > 0 && t != prev) { ranges[prev].y = j; ranges[t].x = j; }
            if (j == R - 1) ranges[t].y = R;
            prev = t;
        }
    }
}

// Tile order for the render kernels: longest list first, to the nearest half power of two (the order only
// shapes the schedule, never the result).  One block: bucket = 2 log2(len) rounded down, histogram, scan,
// scatter -- a few microseconds instead of a general sort.
__global__ void __launch_bounds__(1024) tile_order_kernel(int T, const int2* __restrict__ ranges, uint32_t* __restrict__ order) {
    __shared__ unsigned s_count[64], s_base[64];
    const int t = threadIdx.x;
    if (t < 64) s_count[t] = 0;
    __syncthreads();
    auto bucket = [](unsigned len) -> int {
        if (len == 0) return 63;
        const int lg = 31 - __clz(len);
        const int half = (lg > 0 && ((len >> (lg - 1)) & 1u)) ? 1 : 0;
        return 62 - (2 * lg + half);          // longest lists -> smallest bucket index
    };
    for (int i = t; i < T; i += blockDim.x) {
        const int2 r = ranges[i];
        atomicAdd(&s_count[bucket((unsigned)(r.y - r.x))], 1u);
    }
    __syncthreads();
    if (t == 0) {
        unsigned run = 0;
        for (int b = 0; b < 64; b++) { s_base[b] = run; run += s_count[b]; }
    }
    __syncthreads();
    for (int i = t; i < T; i += blockDim.x) {
        const int2 r = ranges[i];
        order[atomicAdd(&s_base[bucket((unsigned)(r.y - r.x))], 1u)] = (uint32_t)i;
    }
}

template <typename KeyT>
__global__ void export_keys_kernel(int R, const KeyT* __restrict__ tile_keys, const uint32_t* __restrict__ vals,
                                   const float* __restrict__ depths, uint64_t* __restrict__ out) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= R) return;
    out[j] = ((uint64_t)tile_keys[j] << 32) | (uint64_t)__float_as_uint(depths[vals[j]]);
}

__global__ void export_tiles_kernel(int P, const uint2* __restrict__ rect, uint32_t* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint32_t wh = rect[i].y;
    out[i] = (wh & 0xffffu) * (wh >> 16);
}

}  // namespace

cudaError_t launch_depth_order(GeomView gv, int P, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    size_t bytes = gv.dsort_tmp_bytes;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(gv.dsort_tmp, bytes, reinterpret_cast<const uint32_t*>(gv.depths),
                                                    gv.depth_sorted, gv.iota, gv.order, P, 0, 32, s);
    if (e != cudaSuccess) return e;
    TilesIt<TilesInDepthOrder> it(cub::CountingInputIterator<uint32_t>(0), TilesInDepthOrder{gv.order, gv.rect});
    bytes = gv.scan_tmp_bytes;
    return cub::DeviceScan::InclusiveSum(gv.scan_tmp, bytes, it, gv.offsets, P, s);
}

cudaError_t launch_emit(GeomView gv, int P, int W, int H, BinView bv, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    const int gx = (W + TILE - 1) / TILE;
    const int blocks = (P + EMIT_THREADS - 1) / EMIT_THREADS;
    if (bv.key_bytes == 2)
        emit_keys_kernel<uint16_t><<<blocks, EMIT_THREADS, 0, s>>>(P, gx, gv.order, gv.rect, gv.offsets,
                                                                  (uint16_t*)bv.keys_unsorted, bv.vals_unsorted);
    else
        emit_keys_kernel<uint32_t><<<blocks, EMIT_THREADS, 0, s>>>(P, gx, gv.order, gv.rect, gv.offsets,
                                                                  (uint32_t*)bv.keys_unsorted, bv.vals_unsorted);
    return cudaGetLastError();
}

cudaError_t launch_tile_sort(BinView bv, int R, int num_tiles, cudaStream_t s) {
    if (R == 0) return cudaSuccess;
    size_t bytes = bv.sort_tmp_bytes;
    const int bits = tile_bits(num_tiles);
    if (bv.key_bytes == 2)
        return cub::DeviceRadixSort::SortPairs(bv.sort_tmp, bytes, (const uint16_t*)bv.keys_unsorted, (uint16_t*)bv.keys_sorted,
                                               bv.vals_unsorted, bv.vals_sorted, R, 0, bits, s);
    return cub::DeviceRadixSort::SortPairs(bv.sort_tmp, bytes, (const uint32_t*)bv.keys_unsorted, (uint32_t*)bv.keys_sorted,
                                           bv.vals_unsorted, bv.vals_sorted, R, 0, bits, s);
}

cudaError_t launch_ranges(BinView bv, int R, ImgView iv, int W, int H, cudaStream_t s) {
    const size_t T = (size_t)tiles_of(W, H);
    if (T == 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(iv.ranges, 0, T * sizeof(int2), s);
    if (e != cudaSuccess) return e;
    if (R > 0) {
        if (bv.key_bytes == 2)
            identify_tile_ranges_kernel<uint16_t><<<(R + 256 * 8 - 1) / (256 * 8), 256, 0, s>>>(R, (const uint16_t*)bv.keys_sorted, iv.ranges);
        else
            identify_tile_ranges_kernel<uint32_t><<<(R + 256 * 4 - 1) / (256 * 4), 256, 0, s>>>(R, (const uint32_t*)bv.keys_sorted, iv.ranges);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    // longest list first: the render kernels' critical path is their heaviest tile
    tile_order_kernel<<<1, 1024, 0, s>>>((int)T, iv.ranges, iv.tile_order);
    return cudaGetLastError();
}

// Parity / inspection only: rebuild the 64-bit keys, tiles touched and the index-order offsets.
cudaError_t launch_export_binning(int P, int R, GeomView gv, BinView bv, const b200splat_state_out& o, cudaStream_t s) {
    if (o.keys_sorted && R > 0) {
        if (bv.key_bytes == 2)
            export_keys_kernel<uint16_t><<<(R + 255) / 256, 256, 0, s>>>(R, (const uint16_t*)bv.keys_sorted, bv.vals_sorted, gv.depths, o.keys_sorted);
        else
            export_keys_kernel<uint32_t><<<(R + 255) / 256, 256, 0, s>>>(R, (const uint32_t*)bv.keys_sorted, bv.vals_sorted, gv.depths, o.keys_sorted);
    }
    if (o.tiles_touched && P > 0) export_tiles_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, gv.rect, o.tiles_touched);
    if (o.point_offsets && P > 0) {
        // the forward's scan scratch is dead once the forward has finished
        TilesIt<TilesInIndexOrder> it(cub::CountingInputIterator<uint32_t>(0), TilesInIndexOrder{gv.rect});
        size_t bytes = gv.scan_tmp_bytes;
        cudaError_t e = cub::DeviceScan::InclusiveSum(gv.scan_tmp, bytes, it, o.point_offsets, P, s);
        if (e != cudaSuccess) return e;
    }
    return cudaGetLastError();
}

}  // namespace b200splat
