// Device-side BVH build (LBVH) for scenes that are created, rendered once and destroyed — the call pattern of the
// reference's Image::par_render (src/image.rs:65-81: every frame starts from a World).  The host's binned-SAH build
// costs ~1.0-1.9 ms for the teapot's 6 320 triangles (more when the ranks of a multi-GPU run share the host cores);
// here the GPU, idle during scene creation, does it in a handful of small launches:
//
//   morton_keys     item -> 30-bit Morton code of its box centre inside the root box; key = code << 32 | item index
//                   (unique keys: no duplicate handling in the hierarchy)
//   sort_keys       up to 8 192 keys: one block, bitonic sort in shared memory;
//   radix_*         more: a stable least-significant-digit radix sort over the 30-bit codes, 4 bits per pass: per pass a
//                   per-tile digit histogram (ballot + popc per digit), an exclusive scan over (digit, tile) and a stable
//                   scatter — "stable" makes the result equal to sorting the unique keys code << 32 | item
//   build_hierarchy Karras 2012: internal node i covers a range of sorted keys found by binary searches on the common
//                   prefix length; children / parents / ranges written in one pass
//   fit_boxes       bottom-up: each leaf walks towards the root, the second arrival at a node (atomic counter) unites
//                   the children's boxes
//   number_nodes    subtrees of <= leaf_max items collapse into leaves (the traversal's leaf format); the surviving internal
//                   nodes are numbered by a block-wide prefix sum
//   emit_nodes      ... and written as 64-byte BvhNodes (both child boxes + links), exactly what the SAH build produces
//                   for the traversal kernels
//   permute_items   leaf items (header + triangle record) gathered into sorted order, so a leaf's items are contiguous
//
// Culling structure only: the boxes are the same padded, outward-rounded fp32 item boxes the host build uses, so the
// walk stays conservative and the rendered frame is bit-identical to the SAH build's and to brute force (tests).
// The tree is of lower quality than the SAH tree (more node visits per ray); rtc_scene_desc.build_flags selects it.
#include <cuda_runtime.h>
#include <stdint.h>

#include "render_kernels.h"

namespace rtcdev {

namespace {

__device__ __forceinline__ uint32_t expand_bits(uint32_t v) {  // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void morton_keys(LbvhArgs a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const float4 lo = a.box[2 * i], hi = a.box[2 * i + 1];
    const float c[3] = {0.5f * lo.x + 0.5f * hi.x, 0.5f * lo.y + 0.5f * hi.y, 0.5f * lo.z + 0.5f * hi.z};
    uint32_t q[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float u = (c[k] - a.root_lo[k]) * a.root_inv_extent[k];
        u = fminf(fmaxf(u, 0.0f), 1.0f);  // (NaN -> 0: a degenerate axis has inv_extent 0)
        q[k] = (uint32_t)fminf(u * 1024.0f, 1023.0f);
    }
    const uint32_t code = expand_bits(q[0]) << 2 | expand_bits(q[1]) << 1 | expand_bits(q[2]);
    a.keys[i] = (unsigned long long)code << 32 | (uint32_t)i;
}

__global__ void __launch_bounds__(1024) sort_keys(LbvhArgs a, int np2) {
    extern __shared__ unsigned long long sk[];
    for (int i = threadIdx.x; i < np2; i += blockDim.x) sk[i] = i < a.n ? a.keys[i] : ~0ull;
    for (int k = 2; k <= np2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            __syncthreads();
            for (int t = threadIdx.x; t < (np2 >> 1); t += blockDim.x) {  // one compare-exchange per pair (i, i + j)
                const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                const int p = i | j;
                const unsigned long long x = sk[i], y = sk[p];
                if ((x > y) == ((i & k) == 0)) { sk[i] = y; sk[p] = x; }
            }
        }
    __syncthreads();
    for (int i = threadIdx.x; i < a.n; i += blockDim.x) a.keys[i] = sk[i];
}

// ---- multi-block radix sort (n > LBVH_ONE_BLOCK_ITEMS) --------------------------------------------------------
// Keys are sorted by their 30-bit Morton code with a STABLE sort; the items start in index order, so ties end up in
// index order = the order of the unique 64-bit keys.  One tile = 256 consecutive keys = one block; a pass handles one
// 4-bit digit:
//   radix_count    hist[digit * tiles + tile] = keys of the tile with that digit
//   radix_scan     exclusive scan of hist in (digit, tile) order: where each tile's keys of each digit start
//   radix_scatter  every key moves to offset[digit][tile] + (its rank among the tile's keys of that digit, in tile order)
constexpr int RADIX_TILE = 256, RADIX_BITS = 4, RADIX_BINS = 1 << RADIX_BITS;

// per-thread digit -> (rank of this key among the tile's keys with the same digit that precede it, tile count per digit)
__device__ __forceinline__ void radix_rank(uint32_t digit, bool valid, uint32_t* s_warp /*[8][16]*/, uint32_t& rank, uint32_t& before_me_warps) {
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t my_rank = 0;
#pragma unroll
    for (int b = 0; b < RADIX_BINS; ++b) {
        const unsigned m = __ballot_sync(0xffffffffu, valid && digit == (uint32_t)b);
        if (lane == 0) s_warp[warp * RADIX_BINS + b] = __popc(m);
        if (valid && digit == (uint32_t)b) my_rank = __popc(m & ((1u << lane) - 1u));
    }
    __syncthreads();
    uint32_t before = 0;
    for (unsigned w = 0; w < warp; ++w) before += s_warp[w * RADIX_BINS + (valid ? digit : 0)];
    rank = my_rank;
    before_me_warps = before;
}

__global__ void __launch_bounds__(RADIX_TILE) radix_count(const unsigned long long* __restrict__ keys, int n, int shift, int tiles, uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_warp[(RADIX_TILE / 32) * RADIX_BINS];
    const int i = blockIdx.x * RADIX_TILE + threadIdx.x;
    const bool valid = i < n;
    const uint32_t digit = valid ? (uint32_t)(keys[i] >> shift) & (RADIX_BINS - 1) : 0u;
    uint32_t rank, before;
    radix_rank(digit, valid, s_warp, rank, before);
    if (threadIdx.x < RADIX_BINS) {
        uint32_t total = 0;
        for (int w = 0; w < RADIX_TILE / 32; ++w) total += s_warp[w * RADIX_BINS + threadIdx.x];
        hist[threadIdx.x * tiles + blockIdx.x] = total;
    }
}

// One block: exclusive scan of `count` values in place, 1 024 per step.
__global__ void __launch_bounds__(1024) radix_scan(uint32_t* __restrict__ v, int count) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_base;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    for (int tile = 0; tile < count; tile += 1024) {
        const int i = tile + (int)threadIdx.x;
        const uint32_t x = i < count ? v[i] : 0u;
        uint32_t incl = x;
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, incl, off);
            if ((int)lane >= off) incl += u;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = s_warp[lane];
            for (int off = 1; off < 32; off <<= 1) {
                const uint32_t u = __shfl_up_sync(0xffffffffu, w, off);
                if ((int)lane >= off) w += u;
            }
            s_warp[lane] = w;  // inclusive over warps
        }
        __syncthreads();
        const uint32_t base = s_base + (warp ? s_warp[warp - 1] : 0u);
        if (i < count) v[i] = base + incl - x;
        __syncthreads();
        if (threadIdx.x == 0) s_base += s_warp[31];
        __syncthreads();
    }
}

__global__ void __launch_bounds__(RADIX_TILE) radix_scatter(const unsigned long long* __restrict__ in, unsigned long long* __restrict__ out, int n, int shift, int tiles,
                                                            const uint32_t* __restrict__ offset) {
    __shared__ uint32_t s_warp[(RADIX_TILE / 32) * RADIX_BINS];
    const int i = blockIdx.x * RADIX_TILE + threadIdx.x;
    const bool valid = i < n;
    const unsigned long long key = valid ? in[i] : 0ull;
    const uint32_t digit = valid ? (uint32_t)(key >> shift) & (RADIX_BINS - 1) : 0u;
    uint32_t rank, before;
    radix_rank(digit, valid, s_warp, rank, before);
    if (valid) out[offset[digit * tiles + blockIdx.x] + before + rank] = key;
}

// length of the common prefix of keys i and j; -1 outside the array (Karras 2012, section 3)
__device__ __forceinline__ int delta(const unsigned long long* __restrict__ keys, int n, unsigned long long ki, int j) {
    if (j < 0 || j >= n) return -1;
    return __clzll((long long)(ki ^ keys[j]));
}

__global__ void build_hierarchy(LbvhArgs a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = a.n;
    if (i >= n - 1) return;
    const unsigned long long* keys = a.keys;
    const unsigned long long ki = keys[i];
    const int d = delta(keys, n, ki, i + 1) - delta(keys, n, ki, i - 1) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, ki, i - d);
    int lmax = 2;
    while (delta(keys, n, ki, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, ki, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, ki, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, ki, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + (d < 0 ? d : 0);
    const int first = i < j ? i : j, last = i < j ? j : i;
    // children: internal node k -> k, sorted leaf p -> ~p.  parent[]: [0, n-1) internal nodes, [n, 2n) leaves.
    int left, right;
    if (first == gamma) { left = ~gamma; a.parent[n + gamma] = i; } else { left = gamma; a.parent[gamma] = i; }
    if (last == gamma + 1) { right = ~(gamma + 1); a.parent[n + gamma + 1] = i; } else { right = gamma + 1; a.parent[gamma + 1] = i; }
    a.child[i] = make_int2(left, right);
    a.range[i] = make_int2(first, last);
    if (i == 0) a.parent[0] = -1;
}

__device__ __forceinline__ void child_box(const LbvhArgs& a, int c, float4& lo, float4& hi) {
    if (c < 0) {
        const uint32_t item = (uint32_t)(a.keys[~c] & 0xffffffffu);
        lo = a.box[2 * item];
        hi = a.box[2 * item + 1];
    } else {  // written by another thread before it bumped this node's counter: read past L1
        lo = __ldcg(a.nbox + 2 * c);
        hi = __ldcg(a.nbox + 2 * c + 1);
    }
}

__global__ void fit_boxes(LbvhArgs a) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = a.n;
    if (p >= n) return;
    int depth = 1;
    for (int node = a.parent[n + p]; node > 0; node = a.parent[node]) ++depth;
    atomicMax(a.result + 1, depth);
    int node = a.parent[n + p];
    while (node >= 0) {
        if (atomicAdd(a.counter + node, 1) == 0) return;  // the sibling subtree is not finished: its last thread continues
        __threadfence();
        const int2 ch = a.child[node];
        float4 l0, h0, l1, h1;
        child_box(a, ch.x, l0, h0);
        child_box(a, ch.y, l1, h1);
        a.nbox[2 * node] = make_float4(fminf(l0.x, l1.x), fminf(l0.y, l1.y), fminf(l0.z, l1.z), 0.0f);
        a.nbox[2 * node + 1] = make_float4(fmaxf(h0.x, h1.x), fmaxf(h0.y, h1.y), fmaxf(h0.z, h1.z), 0.0f);
        __threadfence();
        node = a.parent[node];
    }
}

// One block: numbers the internal nodes that survive the collapse (range of more than leaf_max items), in index order,
// 1 024 nodes per step: ballot + popc inside a warp, a shuffle scan over the 32 warp totals, a running base.
__global__ void __launch_bounds__(1024) number_nodes(LbvhArgs a) {
    __shared__ int s_warp[32];
    __shared__ int s_base;
    const int n_int = a.n - 1;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    for (int tile = 0; tile < n_int; tile += 1024) {
        const int i = tile + (int)threadIdx.x;
        bool kept = false;
        if (i < n_int) {
            const int2 r = a.range[i];
            kept = r.y - r.x + 1 > a.leaf_max;
        }
        const unsigned m = __ballot_sync(0xffffffffu, kept);
        if (lane == 0) s_warp[warp] = __popc(m);
        __syncthreads();
        if (warp == 0) {
            int v = s_warp[lane];
            for (int off = 1; off < 32; off <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, v, off);
                if ((int)lane >= off) v += u;
            }
            s_warp[lane] = v;  // inclusive
        }
        __syncthreads();
        const int base = s_base + (warp ? s_warp[warp - 1] : 0);
        if (i < n_int) a.dense[i] = kept ? base + __popc(m & ((1u << lane) - 1u)) : -1;
        __syncthreads();
        if (threadIdx.x == 0) s_base += s_warp[31];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        a.result[2] = s_base;  // number of nodes
        a.result[0] = 0;       // the root is internal node 0, kept (n > leaf_max) and numbered first
    }
}

// Surviving internal node -> 64-byte BvhNode: both child boxes + links (inner node: its number; collapsed subtree or
// single item: leaf link ~(first << 2 | count - 1)).
__global__ void emit_nodes(LbvhArgs a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n - 1) return;
    const int me = a.dense[i];
    if (me < 0) return;
    const int2 ch = a.child[i];
    BvhNode nd;
    int links[2];
    float4 lo[2], hi[2];
    const int c[2] = {ch.x, ch.y};
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        child_box(a, c[k], lo[k], hi[k]);
        if (c[k] < 0) links[k] = ~((~c[k]) << 2);  // one item
        else {
            const int2 r = a.range[c[k]];
            const int count = r.y - r.x + 1;
            links[k] = count > a.leaf_max ? a.dense[c[k]] : ~((r.x << 2) | (count - 1));
        }
    }
    nd.lo0[0] = lo[0].x; nd.lo0[1] = lo[0].y; nd.lo0[2] = lo[0].z;
    nd.hi0[0] = hi[0].x; nd.hi0[1] = hi[0].y; nd.hi0[2] = hi[0].z;
    nd.lo1[0] = lo[1].x; nd.lo1[1] = lo[1].y; nd.lo1[2] = lo[1].z;
    nd.hi1[0] = hi[1].x; nd.hi1[1] = hi[1].y; nd.hi1[2] = hi[1].z;
    nd.c0 = links[0];
    nd.c1 = links[1];
    nd.pad0 = 0;
    nd.pad1 = 0;
    a.bvh[me] = nd;
}

// Leaf records in sorted order, so the items of a leaf are contiguous: the 16-byte header the traversal reads instead of
// prims[], and for triangles the (p1, e1, e2) record next to it.
__global__ void permute_items(LbvhArgs a) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.n) return;
    const uint32_t item = (uint32_t)(a.keys[p] & 0xffffffffu);
    const int32_t id = a.item_ids[item];
    LeafItem li;
    li.id = id;
    li.kind_flags = 0;
    li.xform = 0;
    li.geom = -1;
    double2* d = reinterpret_cast<double2*>(a.leaf_tri + 10 * (size_t)p);
    bool triangle = false;
    if (id >= 0) {
        const Prim pr = a.prims[id];
        li.kind_flags = pr.kind_flags;
        li.xform = pr.xform;
        li.geom = pr.geom;
        triangle = (pr.kind_flags & 0xff) >= 5;
        if (triangle) {
            const double2* s = reinterpret_cast<const double2*>(a.tri_pe + 10 * (size_t)pr.geom);
#pragma unroll
            for (int k = 0; k < 5; ++k) d[k] = s[k];
        }
    }
    if (!triangle) {
#pragma unroll
        for (int k = 0; k < 5; ++k) d[k] = make_double2(0.0, 0.0);
    }
    a.leaf_items[p] = li;
}

}  // namespace

// Enqueues the build on `st`.  a.result (3 ints, zeroed here): [0] root link, [1] depth of the radix tree (an upper
// bound of the collapsed tree's), [2] node count.  Returns a CUDA error code.
int build_lbvh(const LbvhArgs& a, cudaStream_t st) {
    if (a.n < 8 || a.n > LBVH_MAX_ITEMS) return (int)cudaErrorInvalidValue;
    cudaMemsetAsync(a.counter, 0, sizeof(int) * (size_t)a.n, st);
    cudaMemsetAsync(a.result, 0, sizeof(int) * 4, st);
    const int block = 128, grid = (a.n + block - 1) / block;
    morton_keys<<<grid, block, 0, st>>>(a);
    if (a.n <= LBVH_ONE_BLOCK_ITEMS) {
        int np2 = 1;
        while (np2 < a.n) np2 <<= 1;
        const size_t smem = (size_t)np2 * sizeof(unsigned long long);
        // (per device and cheap: set on every build rather than tracking which devices have seen it)
        if (smem > 48u * 1024u &&
            cudaFuncSetAttribute(sort_keys, cudaFuncAttributeMaxDynamicSharedMemorySize, LBVH_ONE_BLOCK_ITEMS * (int)sizeof(unsigned long long)) != cudaSuccess)
            return (int)cudaGetLastError();
        sort_keys<<<1, 1024, smem, st>>>(a, np2);
    } else {
        // 30-bit codes sit in bits 32..61 of the keys: 8 passes of 4 bits, ping-pong between keys and keys_alt (even count: ends in keys)
        const int tiles = (a.n + RADIX_TILE - 1) / RADIX_TILE;
        unsigned long long *src = a.keys, *dst = a.keys_alt;
        for (int shift = 32; shift < 62; shift += RADIX_BITS) {
            radix_count<<<tiles, RADIX_TILE, 0, st>>>(src, a.n, shift, tiles, a.radix_hist);
            radix_scan<<<1, 1024, 0, st>>>(a.radix_hist, RADIX_BINS * tiles);
            radix_scatter<<<tiles, RADIX_TILE, 0, st>>>(src, dst, a.n, shift, tiles, a.radix_hist);
            unsigned long long* t = src; src = dst; dst = t;
        }
        static_assert(((62 - 32 + RADIX_BITS - 1) / RADIX_BITS) % 2 == 0, "an even number of passes leaves the sorted keys in a.keys");
    }
    build_hierarchy<<<grid, block, 0, st>>>(a);
    fit_boxes<<<grid, block, 0, st>>>(a);
    number_nodes<<<1, 1024, 0, st>>>(a);
    emit_nodes<<<grid, block, 0, st>>>(a);
    permute_items<<<grid, block, 0, st>>>(a);
    return (int)cudaGetLastError();
}

}  // namespace rtcdev
