// slb_fast.cuh — speedups.advance_board for boards with 4 <= H, W <= 32 in the reference's uint16
// layout: one warp per board, cells -> bit-planes in registers on the way in, k steps on chip,
// planes -> cells on the way out (one HBM round trip for k steps).
#pragma once
#include "slb_resident.cuh"

namespace slb {

constexpr int FAST_WARPS = 4;                 // warps (= boards) per CTA
constexpr int FAST_STAGE_U16 = 32 * 34;       // per-warp staging area (H <= 32 rows, stride <= 34)

// all boards 8-byte aligned and the shared image equals the memory image
__host__ __device__ inline bool fast_vec_ok(int h, int w, const void *a, const void *b) {
    return row_stride(w) == w && ((size_t)h * w * 2) % 8 == 0 && ((uintptr_t)a % 8) == 0 &&
           ((uintptr_t)b % 8) == 0;
}

// ---------------------------------------------------------------------------------------------
// A1: advance_board, k steps on chip.  grid = ceil(N / FAST_WARPS).
// ---------------------------------------------------------------------------------------------
template <int HW>  // HW = 26: the benchmark board size at compile time; 0: runtime shape
__global__ void __launch_bounds__(FAST_WARPS * 32)
advance_fast_kernel(const uint16_t *__restrict__ in, uint16_t *__restrict__ out, int n, int h_, int w_,
                    const float *__restrict__ spawn_prob, uint64_t seed, uint64_t env_id0,
                    uint64_t *counters, int k_steps, bool vec) {
    __shared__ __align__(16) uint16_t stage[FAST_WARPS][FAST_STAGE_U16];
    const int warp = threadIdx.x >> 5, i = blockIdx.x * FAST_WARPS + warp;
    if (i >= n) return;
    const int h = HW ? HW : h_, w = HW ? HW : w_;
    const Geo g = make_geo(h, w);
    const size_t cells = (size_t)h * w;
    uint32_t P[16];
    load_planes<HW>(in + i * cells, stage[warp], h, w, vec, P);
    uint64_t counter = counters ? counters[i] : 0;
    const float p = spawn_prob[i];
    uint32_t touched = 0;  // this lane's cells that flipped (single step), or all ones once anything changed
    if (k_steps == 1) {
        PhysDelta dl;
        bool quiet;
        counter = physics_compute(P, g, p, seed, env_id0 + i, counter, dl, quiet);
        if (__any_sync(FULL, dl.flip != 0)) {
            physics_apply(P, dl);
            touched = dl.flip;
        }
    } else {
        bool ch, any_ch = false;
        for (int k = 0; k < k_steps; k++) {
            counter = physics_planes(P, g, p, seed, env_id0 + i, counter, ch);
            any_ch |= ch;
        }
        if (any_ch) touched = 0xffffffffu;  // several steps: take the full back-transposition
    }
    const int lane = threadIdx.x & 31;
    const int most = __reduce_max_sync(FULL, __popc(touched));  // busiest row
    if (most > 8) {
        store_planes<HW>(out + i * cells, stage[warp], h, w, vec, P);
    } else {
        // Few cells changed (or none): the staging area still holds the input image.  A cell that flipped
        // lost every bit and, where it was born, got alive / destructible / colour bits only (physics_apply),
        // so its final value is read off five planes and patched into the image, which is then copied out —
        // no back-transposition of the planes.
        const int S = row_stride(w);
        uint16_t *sm = stage[warp];
        uint32_t t = touched;
        while (t) {
            const int x = __ffs(t) - 1;
            t &= t - 1;
            const uint32_t v = ((P[0] >> x) & 1u) | (((P[3] >> x) & 1u) << 3) | (((P[9] >> x) & 1u) << 9) |
                               (((P[10] >> x) & 1u) << 10) | (((P[11] >> x) & 1u) << 11);
            sm[lane * S + x] = (uint16_t)v;
        }
        __syncwarp();
        uint16_t *dst = out + i * cells;
        if (vec) {
            const uint2 *s2 = reinterpret_cast<const uint2 *>(sm);
            uint2 *d2 = reinterpret_cast<uint2 *>(dst);
            for (int j = lane; j < ((int)cells >> 2); j += 32) d2[j] = s2[j];
        } else {
            for (int e = lane; e < (int)cells; e += 32) {
                int r = e / w, c = e - r * w;
                dst[e] = sm[r * S + c];
            }
        }
    }
    if (counters && (threadIdx.x & 31) == 0) counters[i] = counter;
}

// ---------------------------------------------------------------------------------------------
// The same for 26 x 26 boards with BULK ASYNC COPIES (TMA, cp.async.bulk) on both sides.  A board is 1352
// bytes — only 8-byte aligned inside the batch — but two consecutive boards are 2704 = 169 x 16 bytes and
// 16-byte aligned, which is what a bulk copy needs.  Persistent warps: warp v owns pairs v, v + V, ...; one
// lane fetches a pair into the warp's shared staging buffer with ONE instruction (completion on an mbarrier),
// the pair after it is already in flight into the second buffer while this one is stepped (a warp alone keeps
// 2.7 KB of HBM traffic in flight without a register or a load instruction tied up), the boards are stepped
// IN PLACE in the staged image (patching the few flipped cells, or a full back-transposition), and one lane
// sends the image to its place in the output with one bulk store.
// ---------------------------------------------------------------------------------------------
constexpr int PAIR_BYTES = 2 * 26 * 26 * 2;

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void bulk_load_pair(void *dst, const void *src, uint64_t *bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(PAIR_BYTES) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_addr(dst)), "l"(src), "r"(PAIR_BYTES), "r"(smem_addr(bar)) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_addr(bar)), "r"(parity) : "memory");
}

// one advance_board call (k steps) on the board whose cells lie at `sm`, in place
__device__ __forceinline__ uint64_t advance_in_image(uint16_t *sm, float p, uint64_t seed, uint64_t env,
                                                     uint64_t counter, int k_steps) {
    const Geo g = make_geo(26, 26);
    const int lane = threadIdx.x & 31;
    uint32_t P[16];
    planes_from_image<26>(sm, 26, 26, P);
    uint32_t touched = 0;
    if (k_steps == 1) {
        PhysDelta dl;
        bool quiet;
        counter = physics_compute(P, g, p, seed, env, counter, dl, quiet);
        if (__any_sync(FULL, dl.flip != 0)) {
            physics_apply(P, dl);
            touched = dl.flip;
        }
    } else {
        bool ch, any_ch = false;
        for (int k = 0; k < k_steps; k++) {
            counter = physics_planes(P, g, p, seed, env, counter, ch);
            any_ch |= ch;
        }
        if (any_ch) touched = 0xffffffffu;
    }
    const int most = __reduce_max_sync(FULL, __popc(touched));
    __syncwarp();
    if (most > 8) {
        image_from_planes<26>(sm, 26, 26, P);
    } else {
        uint32_t t = touched;  // a flipped cell's final value sits in five planes (see advance_fast_kernel)
        while (t) {
            const int x = __ffs(t) - 1;
            t &= t - 1;
            const uint32_t v = ((P[0] >> x) & 1u) | (((P[3] >> x) & 1u) << 3) | (((P[9] >> x) & 1u) << 9) |
                               (((P[10] >> x) & 1u) << 10) | (((P[11] >> x) & 1u) << 11);
            sm[lane * 26 + x] = (uint16_t)v;
        }
    }
    return counter;
}

__global__ void __launch_bounds__(FAST_WARPS * 32)
advance_pairs_kernel(const uint16_t *in, uint16_t *out, int n_pairs, const float *__restrict__ spawn_prob,
                     uint64_t seed, uint64_t env_id0, uint64_t *counters, int k_steps) {
    __shared__ __align__(128) unsigned char stage[FAST_WARPS][2][PAIR_BYTES];
    __shared__ __align__(8) uint64_t bars[FAST_WARPS][2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int V = gridDim.x * FAST_WARPS;
    int pr = blockIdx.x * FAST_WARPS + warp;
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&bars[warp][0])));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&bars[warp][1])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const unsigned char *src = reinterpret_cast<const unsigned char *>(in);
    unsigned char *dst = reinterpret_cast<unsigned char *>(out);
    if (pr < n_pairs && lane == 0) bulk_load_pair(stage[warp][0], src + (size_t)pr * PAIR_BYTES, &bars[warp][0]);
    int s = 0;
    uint32_t parity0 = 0, parity1 = 0;
    for (; pr < n_pairs; pr += V) {
        // the other buffer was the source of the previous bulk store: its reads must be over before the
        // next pair lands in it
        if (lane == 0) {
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            if (pr + V < n_pairs)
                bulk_load_pair(stage[warp][s ^ 1], src + (size_t)(pr + V) * PAIR_BYTES, &bars[warp][s ^ 1]);
        }
        __syncwarp();
        mbar_wait(&bars[warp][s], s ? parity1 : parity0);
        if (s) parity1 ^= 1u; else parity0 ^= 1u;
#pragma unroll 1
        for (int b = 0; b < 2; b++) {
            const int i = 2 * pr + b;
            uint64_t counter = counters ? counters[i] : 0;
            counter = advance_in_image(reinterpret_cast<uint16_t *>(stage[warp][s] + b * (PAIR_BYTES / 2)),
                                       spawn_prob[i], seed, env_id0 + (uint64_t)i, counter, k_steps);
            if (counters && lane == 0) counters[i] = counter;
        }
        // the image was written through the generic proxy: make it visible to the bulk (async proxy) store
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                         ::"l"(dst + (size_t)pr * PAIR_BYTES), "r"(smem_addr(stage[warp][s])), "r"(PAIR_BYTES) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        s ^= 1;
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// set around a launch whose boards lie in mapped HOST memory (slb_advance_board_host's zero-copy path): those
// take the load / store kernel
static thread_local bool t_advance_no_bulk = false;

inline void launch_advance_fast(const uint16_t *in, uint16_t *out, int n, int h, int w, const float *p,
                                uint64_t seed, uint64_t env0, uint64_t *ctr, int k, cudaStream_t st) {
    if (h == 26 && w == 26) {
        // pairs of boards through bulk async copies (both arrays 16-byte aligned; SLB_ADVANCE_BULK=0: off);
        // an odd last board takes the one-board-per-warp kernel.  Only for ONE step per call, the HBM-bound
        // case: measured on B200 (round 2) still boards 0.122 -> 0.068 ms per 131 072 boards (0.44 -> 0.79 of
        // HBM), spawner boards 0.182 -> 0.168 ms; with k = 32 steps on chip the call is bound by the integer
        // pipe and the persistent warps' static share of boards costs 15-22 % against one board per warp
        static const bool bulk = []() { const char *v = getenv("SLB_ADVANCE_BULK"); return !v || v[0] != '0'; }();
        int done = 0;
        if (bulk && k == 1 && !t_advance_no_bulk && n >= 2 && ((uintptr_t)in % 16) == 0 && ((uintptr_t)out % 16) == 0) {
            static const int max_ctas = []() {
                int dev = 0, sms = 148, per = 1;
                cudaGetDevice(&dev);
                cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
                cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, advance_pairs_kernel, FAST_WARPS * 32, 0);
                return sms * (per > 0 ? per : 1);
            }();
            const int pairs = n / 2;
            int grid = (pairs + FAST_WARPS - 1) / FAST_WARPS;
            if (grid > max_ctas) grid = max_ctas;
            advance_pairs_kernel<<<grid, FAST_WARPS * 32, 0, st>>>(in, out, pairs, p, seed, env0, ctr, k);
            done = 2 * pairs;
        }
        if (done < n) {
            const size_t off = (size_t)done * 26 * 26;
            advance_fast_kernel<26><<<(n - done + FAST_WARPS - 1) / FAST_WARPS, FAST_WARPS * 32, 0, st>>>(
                in + off, out + off, n - done, h, w, p + done, seed, env0 + done, ctr ? ctr + done : nullptr, k,
                fast_vec_ok(h, w, in + off, out + off));
        }
        return;
    }
    int grid = (n + FAST_WARPS - 1) / FAST_WARPS;
    advance_fast_kernel<0><<<grid, FAST_WARPS * 32, 0, st>>>(in, out, n, h, w, p, seed, env0, ctr, k,
                                                             fast_vec_ok(h, w, in, out));
}

}  // namespace slb

#include "slb_fast_side.cuh"
