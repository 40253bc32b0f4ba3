// binning.cu -- tile binning, sm_100a.  HBM / L2-bound integer work; every kernel is hand-written (no library
// sort or scan), and nothing waits on the host.
//
// Stands in for the binning stages of _C.rasterize_gaussians named by BASELINE.json's north_star
// ("duplicateWithKeys pass emitting 64-bit tile|depth keys, then a radix sort and identifyTileRanges").
// The RESULT is bit-identical to that formulation -- every tile's list is ordered by (depth bits,
// Gaussian id), exactly what a stable sort of (tile << 32 | depth) keys emitted in Gaussian order
// yields -- but the traffic is a fraction of it (see splat_internal.cuh "Binning order"):
//
//   1. depth order   stable LSD radix sort of the V visible (depth bits, id) pairs, 4 passes.  The first pass reads
//                    the P depth keys in index order, drops the culled ones and supplies the ids itself (no iota
//                    array, no separate compaction); the digit width adapts to the key range seen by the
//                    preprocess kernel (min / max in the control block): 26 significant bits -> 4 x 7.
//   2. offsets       the last depth pass adds the tiles touched by every Gaussian to the pair count of the emit
//                    block its depth position falls in; one small kernel scans those ~V/256 counts
//   3. emit          block-cooperative duplicateWithKeys: a block owns 256 consecutive Gaussians of the
//                    depth order and therefore a CONTIGUOUS output range (it scans its own 256 counts);
//                    consecutive threads write consecutive (tile id, Gaussian id) slots: fully coalesced
//   4. tile sort     stable LSD sort on the tile id, 2 passes: by tile column (tile % gx), then by tile row
//                    (tile / gx).  The second pass writes only the Gaussian ids (nobody needs sorted keys) and
//                    publishes the tile ranges itself: all pairs of a tile are contiguous in its output, so the
//                    first / last position a block writes for a tile go to ranges[] through atomicMax.
//   5. tile order    longest list first (also turns the atomicMax encoding of the range starts back into starts)
//
// One radix pass = three kernels, none of which can wait on another block (no decoupled look-back, no spinning):
//   count    per block of 4096 elements, a shared-memory digit histogram -> matrix[digit][block]
//   colscan  one warp per digit: exclusive scan of its matrix row over the blocks, row total -> totals[digit]
//   scatter  per block: digit bases (scan of totals) + its matrix column = where its elements of each digit go;
//            stable ranks inside the block from warp-level match_any (a warp owns 512 consecutive elements,
//            16 rounds of 32 in order) + a prefix over the 8 warps; the block is staged in shared memory in
//            sorted order so that every digit leaves as one contiguous run (scattered 4-byte stores cost a
//            whole 32-byte sector each at the L2: measured 89 G sectors/s, 3x the time of the staged form).
// The element counts (V, R) live in the image buffer's control block and are read by the kernels; grids are
// sized for the host-known bounds (P, pair capacity) and surplus blocks exit at once.
#include "splat_internal.cuh"

namespace b200splat {

namespace {

#ifndef RS_MINB
#define RS_MINB 4                      // resident scatter blocks per SM (64 registers, <= 54 KB of shared memory each)
#endif
constexpr unsigned FULLMASK = 0xffffffffu;
constexpr int RS_THREADS = 256, RS_WARPS = RS_THREADS / 32, RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;              // 4096 elements per block
constexpr int DEPTH_MAXBINS = 256;                          // 4 passes x <= 8 bits cover any 32-bit key range
constexpr int DEPTH_PASSES = 4;
constexpr int TILE_MAXBINS = 512;                           // tile columns / rows: images up to 8192 x 8192
constexpr int EMIT_THREADS = 256;                           // Gaussians per block of the emit kernel

// blocks covering n elements, rounded up to a multiple of 4 (the column scan walks matrix rows 16 bytes at a time)
inline int radix_blocks(size_t n) {
    const int b = (int)((n + RS_TILE - 1) / RS_TILE);
    return b > 0 ? (b + 3) & ~3 : 4;
}

// Element count of a pass: a device word (u64 when `wide`) clipped to a host-known bound.
struct Count {
    const unsigned* ptr;
    unsigned bound;
    int wide;
};
__device__ __forceinline__ unsigned load_count(Count c) {
    if (!c.ptr) return c.bound;
    unsigned v = __ldg(c.ptr);
    if (c.wide && __ldg(c.ptr + 1) != 0u) v = 0xffffffffu;
    return min(v, c.bound);
}

// ---- digit extraction ------------------------------------------------------------------------------
struct DepthTraits {
    using KeyT = uint32_t;
    static constexpr int MAXBINS = DEPTH_MAXBINS;
    const unsigned* ctr;      // control block: key range of the visible Gaussians
    int pass;
    const uint2* rect;        // last pass only: tiles touched by the Gaussian landing at depth position `pos` are added to
    uint32_t* emit_part;      //   emit_part[pos / EMIT_THREADS] (the per-block pair counts the emit kernel's offsets come from)
    uint32_t kmin, mask;      // set by init()
    int shift, nbins;
    // (Measured and rejected, round 2: 3 passes of 9 bits when the key range allows.  512 bins per 4096-element block
    // make the per-block histogram / cursor work as large as the ranking itself: 0.182 ms against 0.167 ms for 4 x 7.)
    __device__ __forceinline__ void init() {
        const uint32_t kmn = ~__ldg(ctr + CTR_KMIN_NEG), kmx = __ldg(ctr + CTR_KMAX);
        const uint32_t span = kmx >= kmn ? kmx - kmn : 0u;
        const int bits = 32 - __clz((int)span);             // __clz(0) = 32
        const int w = max(1, (bits + DEPTH_PASSES - 1) / DEPTH_PASSES);      // the passes cover `bits`; w <= 8
        kmin = kmn; shift = pass * w; mask = (1u << w) - 1u; nbins = 1 << w;
    }
    __device__ __forceinline__ bool valid(uint32_t k) const { return k != 0xffffffffu; }
    __device__ __forceinline__ uint32_t digit(uint32_t k) const { return ((k - kmin) >> shift) & mask; }
    __device__ __forceinline__ int bins() const { return nbins; }
    // called by a whole warp (inactive lanes pass live = false); the lanes that land in the same emit block -- nearly
    // always all of them: they write consecutive positions -- add their tiles with ONE atomic
    __device__ __forceinline__ void on_output(bool live, uint32_t pos, uint32_t id) const {
        if (!emit_part) return;
        uint32_t tiles = 0;
        if (live) {
            const uint32_t wh = __ldg(&rect[id].y);
            tiles = (wh & 0xffffu) * (wh >> 16);
        }
        const uint32_t blk = live ? pos / EMIT_THREADS : 0xffffffffu;
        const unsigned peers = __match_any_sync(FULLMASK, blk);
        const uint32_t sum = __reduce_add_sync(peers, tiles);
        if (live && (peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0u) atomicAdd(emit_part + blk, sum);
    }
};

template <typename K>
struct TileTraits {
    using KeyT = K;
    static constexpr int MAXBINS = TILE_MAXBINS;
    int gx, nbins, by_row;
    uint32_t magic;            // floor(2^32 / gx) + 1: tile / gx = umulhi(tile, magic) exactly for tile < 2^20, 2 <= gx <= 512
    __device__ __forceinline__ void init() {}
    __device__ __forceinline__ bool valid(K) const { return true; }
    __device__ __forceinline__ uint32_t digit(K k) const {
        const uint32_t t = (uint32_t)k;
        const uint32_t q = gx == 1 ? t : __umulhi(t, magic);
        return by_row ? q : t - q * (uint32_t)gx;
    }
    __device__ __forceinline__ int bins() const { return nbins; }
    __device__ __forceinline__ void on_output(bool, uint32_t, uint32_t) const {}
};

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit_wait_all() {
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// In-place exclusive scan of s[0..n), n <= 2048, by a 256-thread block (s fully written and synchronised by the
// caller; synchronised again on return).  Returns the total.
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t* s, int n, uint32_t* s_warp) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int per = (n + RS_THREADS - 1) / RS_THREADS;      // <= 8
    const int b0 = tid * per;
    uint32_t v[8], sum = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        v[k] = (k < per && b0 + k < n) ? s[b0 + k] : 0u;
        sum += v[k];
    }
    uint32_t incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(FULLMASK, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t run = incl - sum, total = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; w++) {
        const uint32_t t = s_warp[w];
        if (w < warp) run += t;
        total += t;
    }
#pragma unroll
    for (int k = 0; k < 8; k++) {
        if (k < per && b0 + k < n) { s[b0 + k] = run; run += v[k]; }
    }
    __syncthreads();
    return total;
}

// ---- one radix pass: count / colscan / scatter ---------------------------------------------------------
// count: 16-byte loads (8 u16 / 4 u32 keys each), all of a thread's loads in flight before the first histogram update
template <class TR>
__global__ void __launch_bounds__(RS_THREADS) radix_count_kernel(TR tr, const typename TR::KeyT* __restrict__ keys, Count cnt,
                                                                 int stride, uint32_t* __restrict__ matrix,
                                                                 uint32_t* __restrict__ zero_ptr, int zero_n) {
    using KeyT = typename TR::KeyT;
    constexpr int VEC = 16 / sizeof(KeyT), NLOAD = RS_ITEMS / VEC;
    __shared__ unsigned s_hist[TR::MAXBINS];
    // housekeeping for a later stage (the emit kernel's per-block pair counts start from zero)
    for (int i = blockIdx.x * RS_THREADS + threadIdx.x; i < zero_n; i += gridDim.x * RS_THREADS) zero_ptr[i] = 0u;
    const unsigned n = load_count(cnt);
    const unsigned base = blockIdx.x * (unsigned)RS_TILE;
    if (base >= n) return;
    tr.init();
    const int bins = tr.bins(), tid = threadIdx.x;
    for (int d = tid; d < bins; d += RS_THREADS) s_hist[d] = 0u;
    uint4 raw[NLOAD];
#pragma unroll
    for (int l = 0; l < NLOAD; l++) {
        const unsigned idx = base + (unsigned)(l * RS_THREADS + tid) * VEC;
        raw[l] = idx < n ? __ldg(reinterpret_cast<const uint4*>(keys + idx)) : make_uint4(0u, 0u, 0u, 0u);   // padded arrays: in bounds
    }
    __syncthreads();
#pragma unroll
    for (int l = 0; l < NLOAD; l++) {
        const unsigned idx = base + (unsigned)(l * RS_THREADS + tid) * VEC;
        const KeyT* kk = reinterpret_cast<const KeyT*>(&raw[l]);
#pragma unroll
        for (int e = 0; e < VEC; e++)
            if (idx + e < n && tr.valid(kk[e])) atomicAdd(&s_hist[tr.digit(kk[e])], 1u);
    }
    __syncthreads();
    for (int d = tid; d < bins; d += RS_THREADS) matrix[(size_t)d * stride + blockIdx.x] = s_hist[d];
}

// colscan: one block per digit.  Its 8 warps take consecutive eighths of the matrix row (a lane 4 consecutive blocks
// per step: the row stride is a multiple of 4): segment sums first, then every warp scans its segment from its base.
template <class TR>
__global__ void __launch_bounds__(RS_THREADS) radix_colscan_kernel(TR tr, Count cnt, int stride, uint32_t* __restrict__ matrix,
                                                                   uint32_t* __restrict__ totals) {
    __shared__ uint32_t s_seg[RS_WARPS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int d = blockIdx.x;
    tr.init();
    if (d >= tr.bins()) return;
    const unsigned n = load_count(cnt);
    const int nb = (int)((n + RS_TILE - 1) / RS_TILE);
    const int nq = (nb + 3) >> 2;                                         // 16-byte groups in the row
    const int per = ((nq + RS_WARPS - 1) / RS_WARPS + 31) & ~31;          // groups per warp, whole steps of 32
    const int q0 = warp * per, q1 = min(nq, q0 + per);
    uint4* row = reinterpret_cast<uint4*>(matrix + (size_t)d * stride);
    auto load = [&](int q) -> uint4 {
        uint4 v = q < q1 ? row[q] : make_uint4(0u, 0u, 0u, 0u);
        const int b = 4 * q;
        if (b + 1 >= nb) v.y = 0u;
        if (b + 2 >= nb) v.z = 0u;
        if (b + 3 >= nb) v.w = 0u;
        if (b >= nb) v.x = 0u;
        return v;
    };
    uint32_t sum = 0;
    for (int q = q0 + lane; q < q1; q += 32) { const uint4 v = load(q); sum += v.x + v.y + v.z + v.w; }
    sum = __reduce_add_sync(FULLMASK, sum);
    if (lane == 0) s_seg[warp] = sum;
    __syncthreads();
    uint32_t run = 0, total = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; w++) {
        const uint32_t t = s_seg[w];
        if (w < warp) run += t;
        total += t;
    }
    for (int qs = q0; qs < q1; qs += 32) {
        const int q = qs + lane;
        const uint4 v = load(q);
        const uint32_t part = v.x + v.y + v.z + v.w;
        uint32_t incl = part;
#pragma unroll
        for (int sft = 1; sft < 32; sft <<= 1) {
            const uint32_t t = __shfl_up_sync(FULLMASK, incl, sft);
            if (lane >= sft) incl += t;
        }
        const uint32_t e0 = run + incl - part;
        if (q < q1) row[q] = make_uint4(e0, e0 + v.x, e0 + v.x + v.y, e0 + v.x + v.y + v.z);
        run += __shfl_sync(FULLMASK, incl, 31);
    }
    if (threadIdx.x == 0) totals[d] = total;
}

// scatter.  Shared memory (dynamic): digit bases, per-warp digit counts, the block's 4096 (key, value) pairs as they
// arrive (cp.async, in flight while the digit bases are scanned); the sorted order then takes the place of the input.
template <class TR>
struct ScatterSmem {
    using KeyT = typename TR::KeyT;
    static constexpr int MB = TR::MAXBINS;
    static constexpr size_t off_base = 0;
    static constexpr size_t off_lstart = off_base + sizeof(uint32_t) * MB;
    static constexpr size_t off_whist = off_lstart + sizeof(uint32_t) * MB;
    static constexpr size_t off_ikey = off_whist + sizeof(uint32_t) * RS_WARPS * MB;
    static constexpr size_t off_ival = off_ikey + sizeof(KeyT) * RS_TILE;
    static constexpr size_t off_okey = off_ikey;             // the sorted order overwrites the input (through registers):
    static constexpr size_t off_oval = off_ival;             //   42-52 KB per block instead of 74-84, 4 blocks per SM instead of 3
    static constexpr size_t off_warp = off_ival + sizeof(uint32_t) * RS_TILE;
    static constexpr size_t bytes = off_warp + sizeof(uint32_t) * RS_WARPS;
};

template <class TR>
__global__ void __launch_bounds__(RS_THREADS, RS_MINB) radix_scatter_kernel(TR tr, const typename TR::KeyT* __restrict__ keys_in,
                                                                      const uint32_t* __restrict__ vals_in,
                                                                      typename TR::KeyT* __restrict__ keys_out,
                                                                      uint32_t* __restrict__ vals_out, Count cnt, int stride,
                                                                      const uint32_t* __restrict__ matrix,
                                                                      const uint32_t* __restrict__ totals, int2* __restrict__ ranges) {
    using KeyT = typename TR::KeyT;
    using SM = ScatterSmem<TR>;
    constexpr int MB = TR::MAXBINS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* s_base = reinterpret_cast<uint32_t*>(smem_raw + SM::off_base);      // global position of the block's first element of each digit
    uint32_t* s_lstart = reinterpret_cast<uint32_t*>(smem_raw + SM::off_lstart);  // first slot of each digit in the block's sorted order
    uint32_t (*s_whist)[MB] = reinterpret_cast<uint32_t (*)[MB]>(smem_raw + SM::off_whist);   // per-warp counts -> prefix over warps
    KeyT* s_ikey = reinterpret_cast<KeyT*>(smem_raw + SM::off_ikey);
    uint32_t* s_ival = reinterpret_cast<uint32_t*>(smem_raw + SM::off_ival);
    KeyT* s_okey = reinterpret_cast<KeyT*>(smem_raw + SM::off_okey);
    uint32_t* s_oval = reinterpret_cast<uint32_t*>(smem_raw + SM::off_oval);
    uint32_t* s_warp = reinterpret_cast<uint32_t*>(smem_raw + SM::off_warp);

    const unsigned n = load_count(cnt);
    const unsigned base = blockIdx.x * (unsigned)RS_TILE;
    if (base >= n) return;
    tr.init();
    const int m = (int)min((unsigned)RS_TILE, n - base);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // the block's input on its way to shared memory (16-byte chunks; the arrays are padded, a chunk never leaves them)
    {
        constexpr int KV = 16 / sizeof(KeyT);
        const int kchunks = (m + KV - 1) / KV, vchunks = (m + 3) / 4;
        for (int c = tid; c < kchunks; c += RS_THREADS) cp_async16(s_ikey + c * KV, keys_in + base + c * KV);
        if (vals_in)
            for (int c = tid; c < vchunks; c += RS_THREADS) cp_async16(s_ival + c * 4, vals_in + base + c * 4);
    }
    const int bins = tr.bins();
    for (int d = tid; d < bins; d += RS_THREADS) {
        s_base[d] = __ldg(totals + d);
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) s_whist[w][d] = 0;
    }
    __syncthreads();
    block_excl_scan(s_base, bins, s_warp);
    for (int d = tid; d < bins; d += RS_THREADS) s_base[d] += __ldg(matrix + (size_t)d * stride + blockIdx.x);
    cp_async_commit_wait_all();
    __syncthreads();

    // A warp owns 512 consecutive elements: 16 rounds of 32, in order.
    // 1. per-warp digit counts (shared-memory atomics, any order)
    const int t0 = warp * (RS_ITEMS * 32) + lane;
#pragma unroll
    for (int e = 0; e < RS_ITEMS; e++) {
        const int t = t0 + e * 32;
        if (t < m) {
            const KeyT key = s_ikey[t];
            if (tr.valid(key)) atomicAdd(&s_whist[warp][tr.digit(key)], 1u);
        }
    }
    __syncthreads();
    // 2. counts -> cursors: slot of the warp's first element of each digit in the block's sorted order
    for (int d = tid; d < bins; d += RS_THREADS) {
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) {
            const uint32_t c = s_whist[w][d];
            s_whist[w][d] = run;
            run += c;
        }
        s_lstart[d] = run;
    }
    __syncthreads();
    const int mv = (int)block_excl_scan(s_lstart, bins, s_warp);          // elements of the block that take part
    for (int d = tid; d < bins; d += RS_THREADS) {
        const uint32_t ls = s_lstart[d];
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) s_whist[w][d] += ls;
    }
    __syncthreads();
    // 3. stable slots.  Per round, the lowest lane of every digit group advances the warp's cursor of that digit by the
    // group's size and hands the old value to its group.  Nothing in a round waits for the round before: the match
    // instructions of 8 rounds are issued together, then their 8 shared-memory atomics (a warp's atomics on one
    // address execute in issue order, which is what makes the order stable), then the old values are broadcast and
    // the slots are kept in registers until every warp has read its input (the sorted order overwrites it).
    const unsigned lt = (1u << lane) - 1u;
    uint32_t slot2[RS_ITEMS / 2];          // the 16 slots of this thread, two per word; 0xffff = element takes no part
#pragma unroll
    for (int g = 0; g < RS_ITEMS; g += 8) {
        uint32_t dd[8], old[8];
        unsigned peers[8];
        KeyT key[8];
#pragma unroll
        for (int e = 0; e < 8; e++) {
            const int t = t0 + (g + e) * 32;
            key[e] = s_ikey[t < m ? t : 0];
            dd[e] = (t < m && tr.valid(key[e])) ? tr.digit(key[e]) : 0xffffffffu;
            peers[e] = __match_any_sync(FULLMASK, dd[e]);
        }
#pragma unroll
        for (int e = 0; e < 8; e++) {
            old[e] = 0u;
            if (dd[e] != 0xffffffffu && (peers[e] & lt) == 0u) old[e] = atomicAdd(&s_whist[warp][dd[e]], (uint32_t)__popc(peers[e]));
        }
#pragma unroll
        for (int e = 0; e < 8; e++) {
            const uint32_t o = __shfl_sync(FULLMASK, old[e], __ffs(peers[e]) - 1);
            const uint32_t li = dd[e] != 0xffffffffu ? o + __popc(peers[e] & lt) : 0xffffu;
            if (e & 1) slot2[(g + e) >> 1] |= li << 16; else slot2[(g + e) >> 1] = li;
        }
    }
    {   // input -> registers, barrier, registers -> sorted slots of the same arrays
        KeyT kreg[RS_ITEMS];
        uint32_t vreg[RS_ITEMS];
#pragma unroll
        for (int e = 0; e < RS_ITEMS; e++) {
            const int t = t0 + e * 32;
            kreg[e] = s_ikey[t < m ? t : 0];
            vreg[e] = vals_in ? s_ival[t < m ? t : 0] : base + (unsigned)t;
        }
        __syncthreads();
#pragma unroll
        for (int e = 0; e < RS_ITEMS; e++) {
            const uint32_t li = (slot2[e >> 1] >> ((e & 1) * 16)) & 0xffffu;
            if (li != 0xffffu) { s_okey[li] = kreg[e]; s_oval[li] = vreg[e]; }
        }
    }
    __syncthreads();
    for (int d = tid; d < bins; d += RS_THREADS) s_base[d] -= s_lstart[d];      // position = s_base[digit] + slot
    __syncthreads();
    for (int i0 = warp * 32; i0 < mv; i0 += RS_THREADS) {    // warp-uniform trip count: on_output is a warp operation
        const int i = i0 + lane;
        const bool live = i < mv;
        const KeyT key = s_okey[live ? i : 0];
        const uint32_t val = s_oval[live ? i : 0];
        const uint32_t pos = s_base[tr.digit(key)] + (uint32_t)i;
        tr.on_output(live, pos, val);
        if (!live) continue;
        vals_out[pos] = val;
        if (keys_out) keys_out[pos] = key;
        if (ranges) {
            // all pairs of a tile are contiguous in this pass's output: start = min position, end = max position + 1.
            // The start is kept as max(~position) so that zero-filled ranges are the identity (launch_tile_order
            // decodes).  A run boundary inside the block closes the previous tile's run and opens this one's.
            const KeyT prev = i > 0 ? s_okey[i - 1] : key;
            if (i == 0 || prev != key) {
                atomicMax(reinterpret_cast<unsigned*>(&ranges[(uint32_t)key].x), ~pos);
                if (i > 0) atomicMax(reinterpret_cast<unsigned*>(&ranges[(uint32_t)prev].y), s_base[tr.digit(prev)] + (uint32_t)i);
            }
            if (i == mv - 1) atomicMax(reinterpret_cast<unsigned*>(&ranges[(uint32_t)key].y), pos + 1u);
        }
    }
}

template <class TR>
cudaError_t radix_pass(TR tr, const typename TR::KeyT* keys_in, const uint32_t* vals_in, typename TR::KeyT* keys_out,
                       uint32_t* vals_out, Count cnt, int blocks, uint32_t* matrix, uint32_t* totals, int2* ranges,
                       cudaStream_t s, uint32_t* zero_ptr = nullptr, int zero_n = 0) {
    constexpr size_t smem = ScatterSmem<TR>::bytes;
    static bool configured[64] = {false};                    // per instantiation and device: the attribute is per function
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (!configured[dev & 63]) {
        e = cudaFuncSetAttribute(radix_scatter_kernel<TR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured[dev & 63] = true;
    }
    radix_count_kernel<TR><<<blocks, RS_THREADS, 0, s>>>(tr, keys_in, cnt, blocks, matrix, zero_ptr, zero_n);
    radix_colscan_kernel<TR><<<TR::MAXBINS, RS_THREADS, 0, s>>>(tr, cnt, blocks, matrix, totals);
    radix_scatter_kernel<TR><<<blocks, RS_THREADS, smem, s>>>(tr, keys_in, vals_in, keys_out, vals_out, cnt, blocks, matrix, totals, ranges);
    return cudaGetLastError();
}

// ---- inclusive scan (reduce / spine / apply), 4096 elements per block -----------------------------------------
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

template <class F>
__global__ void __launch_bounds__(RS_THREADS) scan_reduce_kernel(F f, Count cnt, uint32_t* __restrict__ part) {
    __shared__ uint32_t s_warp[RS_WARPS];
    const unsigned n = load_count(cnt);
    const unsigned base = blockIdx.x * (unsigned)RS_TILE;
    if (base >= n) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t sum = 0;
#pragma unroll 4
    for (int e = 0; e < RS_ITEMS; e++) {
        const unsigned idx = base + e * RS_THREADS + tid;
        if (idx < n) sum += f(idx);
    }
    sum = __reduce_add_sync(FULLMASK, sum);
    if (lane == 0) s_warp[warp] = sum;
    __syncthreads();
    if (tid == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) t += s_warp[w];
        part[blockIdx.x] = t;
    }
}

// exclusive scan of the block partials, one block (any number of partials = ceil(count / per_block))
__global__ void __launch_bounds__(1024) scan_spine_kernel(Count cnt, int per_block, uint32_t* __restrict__ part) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    const unsigned n = load_count(cnt);
    const int nb = (int)((n + per_block - 1) / per_block);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0u;
    __syncthreads();
    for (int b0 = 0; b0 < nb; b0 += 1024) {
        const int b = b0 + tid;
        const uint32_t v = b < nb ? part[b] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            const uint32_t t = __shfl_up_sync(FULLMASK, incl, s);
            if (lane >= s) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        uint32_t wb = s_carry;
        for (int w = 0; w < warp; w++) wb += s_warp[w];
        if (b < nb) part[b] = wb + incl - v;
        __syncthreads();
        if (tid == 1023) s_carry = wb + incl;
        __syncthreads();
    }
}

template <class F>
__global__ void __launch_bounds__(RS_THREADS) scan_apply_kernel(F f, Count cnt, const uint32_t* __restrict__ part,
                                                                uint32_t* __restrict__ out) {
    __shared__ uint32_t s_warp[RS_WARPS];
    const unsigned n = load_count(cnt);
    const unsigned base = blockIdx.x * (unsigned)RS_TILE;
    if (base >= n) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned wbase = base + warp * (RS_ITEMS * 32);
    uint32_t v[RS_ITEMS];
#pragma unroll
    for (int e = 0; e < RS_ITEMS; e++) {
        const unsigned idx = wbase + e * 32 + lane;
        v[e] = idx < n ? f(idx) : 0u;
    }
    uint32_t run = 0;
#pragma unroll
    for (int e = 0; e < RS_ITEMS; e++) {
        uint32_t incl = v[e];
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            const uint32_t t = __shfl_up_sync(FULLMASK, incl, s);
            if (lane >= s) incl += t;
        }
        v[e] = run + incl;
        run += __shfl_sync(FULLMASK, incl, 31);
    }
    if (lane == 0) s_warp[warp] = run;
    __syncthreads();
    uint32_t wb = __ldg(part + blockIdx.x);
    for (int w = 0; w < warp; w++) wb += s_warp[w];
#pragma unroll
    for (int e = 0; e < RS_ITEMS; e++) {
        const unsigned idx = wbase + e * 32 + lane;
        if (idx < n) out[idx] = wb + v[e];
    }
}

template <class F>
cudaError_t inclusive_scan(F f, Count cnt, int blocks, uint32_t* part, uint32_t* out, cudaStream_t s) {
    scan_reduce_kernel<F><<<blocks, RS_THREADS, 0, s>>>(f, cnt, part);
    scan_spine_kernel<<<1, 1024, 0, s>>>(cnt, RS_TILE, part);
    scan_apply_kernel<F><<<blocks, RS_THREADS, 0, s>>>(f, cnt, part, out);
    return cudaGetLastError();
}

}  // namespace

// ---- buffer layouts -------------------------------------------------------------------------------------
GeomView carve_geom(void* base, int P) {
    GeomView g;
    size_t off = 0;
    char* b = (char*)base;
    size_t n = (size_t)(P > 0 ? P : 1);
    g.depths = (float*)(b + off); off += align256(n * sizeof(float));
    g.rec = (float4*)(b + off); off += align256(n * 3 * sizeof(float4));
    g.rect = (uint2*)(b + off); off += align256(n * sizeof(uint2));
    g.dkey_a = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    g.dval_a = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    g.dkey_b = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    g.dval_b = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    g.order = g.dval_b;                                      // four passes: depths -> a -> b -> a -> b
    g.dblocks = radix_blocks(n);
    g.emit_blocks = (int)((n + EMIT_THREADS - 1) / EMIT_THREADS);
    g.emit_part = (uint32_t*)(b + off); off += align256(((size_t)g.emit_blocks + 1) * sizeof(uint32_t));
    g.scan_part = (uint32_t*)(b + off); off += align256(((size_t)g.dblocks + 1) * sizeof(uint32_t));
    g.dmatrix = (uint32_t*)(b + off); off += align256((size_t)DEPTH_MAXBINS * g.dblocks * sizeof(uint32_t));
    g.dtotals = (uint32_t*)(b + off); off += align256((size_t)DEPTH_MAXBINS * sizeof(uint32_t));
    g.total = off;
    return g;
}

BinView carve_binning(void* base, int cap, int num_tiles) {
    BinView v;
    size_t off = 0;
    char* b = (char*)base;
    size_t n = (size_t)(cap > 0 ? cap : 1);
    v.cap = cap > 0 ? cap : 0;
    v.key_bytes = num_tiles <= 65536 ? 2 : 4;
    v.key_a = (void*)(b + off); off += align256(n * v.key_bytes);
    v.val_a = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    v.key_b = (void*)(b + off); off += align256(n * v.key_bytes);
    v.val_b = (uint32_t*)(b + off); off += align256(n * sizeof(uint32_t));
    v.vals_sorted = v.val_a;
    v.tblocks = radix_blocks(n);
    v.tmatrix = (uint32_t*)(b + off); off += align256((size_t)TILE_MAXBINS * v.tblocks * sizeof(uint32_t));
    v.ttotals = (uint32_t*)(b + off); off += align256((size_t)TILE_MAXBINS * sizeof(uint32_t));
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
    v.counters = (unsigned int*)(b + off); off += align256(CTR_WORDS * sizeof(unsigned int));   // right behind the ranges:
    v.clear_bytes = off;                                                                        //   one clear for both
    v.final_T = (float*)(b + off); off += align256((N > 0 ? N : 1) * sizeof(float));
    v.n_contrib = (uint32_t*)(b + off); off += align256((N > 0 ? N : 1) * sizeof(uint32_t));
    const size_t Tn = T > 0 ? T : 1;
    v.tile_order = (uint32_t*)(b + off); off += align256(Tn * sizeof(uint32_t));
    v.total = off;
    return v;
}

namespace {

// Block-cooperative duplicateWithKeys.  Block b owns depth-order positions [256 b, 256 b + 256) and a CONTIGUOUS range
// of output slots: it starts at part[b], the exclusive prefix of the per-block pair counts (accumulated by the last
// depth pass, scanned by scan_spine_kernel), and the block scans the tiles touched by its own 256 Gaussians itself.
// Slot k of Gaussian g (rect minx, miny, w, h) is tile (miny + k / w) * gx + minx + k % w: row-major over the
// rectangle, as the oracle emits.  V (visible Gaussians) comes from the control block; slots at or beyond the pair
// capacity are not written.
template <typename KeyT>
__global__ void __launch_bounds__(EMIT_THREADS) emit_keys_kernel(int gx, const uint32_t* __restrict__ order,
                                                                 const uint2* __restrict__ rect,
                                                                 const uint32_t* __restrict__ part,
                                                                 KeyT* __restrict__ keys, uint32_t* __restrict__ vals,
                                                                 unsigned cap, unsigned int* __restrict__ counters) {
    __shared__ uint32_t s_end[EMIT_THREADS];     // inclusive offsets of this block's Gaussians
    __shared__ uint32_t s_id[EMIT_THREADS];
    __shared__ uint2 s_rect[EMIT_THREADS];
    __shared__ uint32_t s_wsum[EMIT_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int V = (int)__ldg(counters + CTR_VISIBLE);
    const int j0 = blockIdx.x * EMIT_THREADS;
    if (blockIdx.x == 0 && tid == 0) {           // status of the view (b200splat_forward_status)
        const unsigned lo = __ldg(counters + CTR_PAIRS), hi = __ldg(counters + CTR_PAIRS + 1);
        counters[CTR_STATUS] = lo; counters[CTR_STATUS + 1] = hi;
        counters[CTR_STATUS + 2] = cap;
        counters[CTR_STATUS + 3] = (hi != 0u || lo > cap) ? 1u : 0u;
    }
    if (j0 >= V) return;
    const uint32_t base = __ldg(part + blockIdx.x);
    uint32_t id = 0, tiles = 0;
    uint2 r0 = make_uint2(0u, 0u);
    if (j0 + tid < V) {
        id = __ldg(order + j0 + tid);
        r0 = __ldg(rect + id);
        tiles = (r0.y & 0xffffu) * (r0.y >> 16);
    }
    s_id[tid] = id;
    s_rect[tid] = r0;
    uint32_t incl = tiles;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(FULLMASK, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    uint32_t wb = base;
    for (int w = 0; w < warp; w++) wb += s_wsum[w];
    s_end[tid] = wb + incl;                      // positions past V repeat the last offset (no tiles)
    __syncthreads();
    // Each thread owns four consecutive output slots (aligned to 4 so full groups go out as one vector store
    // per array): one binary search and one division for the first slot, then a linear walk.
    const uint32_t end = min(s_end[EMIT_THREADS - 1], cap);
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
        uint32_t gid = s_id[g];
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
                    r = s_rect[g]; w = r.y & 0xffffu; gend = s_end[g]; gid = s_id[g];
                    col = 0;
                    tile = (r.x >> 16) * (uint32_t)gx + (r.x & 0xffffu);
                }
                kk[e] = (KeyT)tile; vv[e] = gid;
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

// Tile order for the render kernels: longest list first, to the nearest half power of two (the order only
// shapes the schedule, never the result).  One block: bucket = 2 log2(len) rounded down, histogram, scan,
// scatter -- a few microseconds instead of a general sort.  Also finishes the ranges: the second tile pass left
// every start as max(~position) (zero for an empty tile), which becomes the start itself here.
__global__ void __launch_bounds__(1024) tile_order_kernel(int T, int2* __restrict__ ranges, uint32_t* __restrict__ order) {
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
        int2 r = ranges[i];
        r.x = r.y > 0 ? (int)~(unsigned)r.x : 0;
        ranges[i] = r;
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

// parity / inspection: the 64-bit keys (tile << 32 | depth bits) of every sorted pair, rebuilt per tile
__global__ void __launch_bounds__(128) export_keys_kernel(const int2* __restrict__ ranges, const uint32_t* __restrict__ vals,
                                                          const float* __restrict__ depths, uint64_t* __restrict__ out) {
    const int2 r = ranges[blockIdx.x];
    for (int j = r.x + threadIdx.x; j < r.y; j += blockDim.x)
        out[j] = ((uint64_t)blockIdx.x << 32) | (uint64_t)__float_as_uint(depths[vals[j]]);
}

__global__ void export_tiles_kernel(int P, const uint2* __restrict__ rect, uint32_t* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint32_t wh = rect[i].y;
    out[i] = (wh & 0xffffu) * (wh >> 16);
}

__global__ void copy_u32_kernel(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst, Count cnt) {
    const unsigned n = load_count(cnt);
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) dst[i] = src[i];
}

}  // namespace

cudaError_t launch_depth_order(GeomView gv, int P, const unsigned int* counters, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    const int blocks = gv.dblocks;
    const Count all{nullptr, (unsigned)P, 0};                         // first pass: every key, culled ones dropped
    const Count vis{counters + CTR_VISIBLE, (unsigned)P, 0};          // then the V visible
    DepthTraits t0{counters, 0, nullptr, nullptr}, t1{counters, 1, nullptr, nullptr}, t2{counters, 2, nullptr, nullptr},
        t3{counters, 3, gv.rect, gv.emit_part};                       // the last pass also sums the pairs per emit block
    cudaError_t e;
    e = radix_pass(t0, reinterpret_cast<const uint32_t*>(gv.depths), (const uint32_t*)nullptr, gv.dkey_a, gv.dval_a, all, blocks,
                   gv.dmatrix, gv.dtotals, nullptr, s, gv.emit_part, gv.emit_blocks + 1);
    if (e != cudaSuccess) return e;
    e = radix_pass(t1, (const uint32_t*)gv.dkey_a, (const uint32_t*)gv.dval_a, gv.dkey_b, gv.dval_b, vis, blocks, gv.dmatrix,
                   gv.dtotals, nullptr, s);
    if (e != cudaSuccess) return e;
    e = radix_pass(t2, (const uint32_t*)gv.dkey_b, (const uint32_t*)gv.dval_b, gv.dkey_a, gv.dval_a, vis, blocks, gv.dmatrix,
                   gv.dtotals, nullptr, s);
    if (e != cudaSuccess) return e;
    e = radix_pass(t3, (const uint32_t*)gv.dkey_a, (const uint32_t*)gv.dval_a, (uint32_t*)nullptr, gv.order, vis, blocks, gv.dmatrix,
                   gv.dtotals, nullptr, s);
    if (e != cudaSuccess) return e;
    scan_spine_kernel<<<1, 1024, 0, s>>>(vis, EMIT_THREADS, gv.emit_part);     // pair counts per emit block -> their offsets
    return cudaGetLastError();
}

cudaError_t launch_emit(GeomView gv, int P, int W, int H, BinView bv, unsigned int* counters, cudaStream_t s) {
    if (P == 0) return cudaSuccess;
    const int gx = (W + TILE - 1) / TILE;
    const int blocks = (P + EMIT_THREADS - 1) / EMIT_THREADS;
    if (bv.key_bytes == 2)
        emit_keys_kernel<uint16_t><<<blocks, EMIT_THREADS, 0, s>>>(gx, gv.order, gv.rect, gv.emit_part, (uint16_t*)bv.key_a, bv.val_a,
                                                                  (unsigned)bv.cap, counters);
    else
        emit_keys_kernel<uint32_t><<<blocks, EMIT_THREADS, 0, s>>>(gx, gv.order, gv.rect, gv.emit_part, (uint32_t*)bv.key_a, bv.val_a,
                                                                  (unsigned)bv.cap, counters);
    return cudaGetLastError();
}

template <typename K>
static cudaError_t tile_sort_impl(BinView bv, int gx, int gy, ImgView iv, cudaStream_t s) {
    const Count pairs{iv.counters + CTR_PAIRS, (unsigned)bv.cap, 1};
    const uint32_t magic = gx > 1 ? (uint32_t)((1ull << 32) / (unsigned long long)gx + 1ull) : 0u;
    TileTraits<K> by_col{gx, gx, 0, magic}, by_row{gx, gy, 1, magic};
    cudaError_t e = radix_pass(by_col, (const K*)bv.key_a, (const uint32_t*)bv.val_a, (K*)bv.key_b, bv.val_b, pairs, bv.tblocks,
                               bv.tmatrix, bv.ttotals, nullptr, s);
    if (e != cudaSuccess) return e;
    return radix_pass(by_row, (const K*)bv.key_b, (const uint32_t*)bv.val_b, (K*)nullptr, bv.vals_sorted, pairs, bv.tblocks,
                      bv.tmatrix, bv.ttotals, iv.ranges, s);
}

cudaError_t launch_tile_sort(BinView bv, int W, int H, ImgView iv, cudaStream_t s) {
    if (bv.cap == 0) return cudaSuccess;
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    return bv.key_bytes == 2 ? tile_sort_impl<uint16_t>(bv, gx, gy, iv, s) : tile_sort_impl<uint32_t>(bv, gx, gy, iv, s);
}

int max_tile_bins() { return TILE_MAXBINS; }

cudaError_t launch_tile_order(ImgView iv, int W, int H, cudaStream_t s) {
    const int T = tiles_of(W, H);
    if (T == 0) return cudaSuccess;
    // longest list first: the render kernels' critical path is their heaviest tile
    tile_order_kernel<<<1, 1024, 0, s>>>(T, iv.ranges, iv.tile_order);
    return cudaGetLastError();
}

// Parity / inspection only: rebuild the 64-bit keys, tiles touched and the index-order offsets.
cudaError_t launch_export_binning(int P, int W, int H, GeomView gv, BinView bv, ImgView iv, const b200splat_state_out& o, cudaStream_t s) {
    const int T = tiles_of(W, H);
    if (o.keys_sorted && bv.cap > 0 && T > 0) export_keys_kernel<<<T, 128, 0, s>>>(iv.ranges, bv.vals_sorted, gv.depths, o.keys_sorted);
    if (o.vals_sorted && bv.cap > 0)
        copy_u32_kernel<<<296, 256, 0, s>>>(bv.vals_sorted, o.vals_sorted, Count{iv.counters + CTR_PAIRS, (unsigned)bv.cap, 1});
    if (o.tiles_touched && P > 0) export_tiles_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, gv.rect, o.tiles_touched);
    if (o.point_offsets && P > 0) {
        // the forward's scan scratch is dead once the forward has finished
        cudaError_t e = inclusive_scan(TilesInIndexOrder{gv.rect}, Count{nullptr, (unsigned)P, 0}, gv.dblocks, gv.scan_part, o.point_offsets, s);
        if (e != cudaSuccess) return e;
    }
    return cudaGetLastError();
}

}  // namespace b200splat
