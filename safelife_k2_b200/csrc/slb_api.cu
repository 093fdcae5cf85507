// slb_api.cu — the C ABI of libsafelife_b200.so (include/safelife_b200.h): argument checks, kernel
// selection (bit-sliced fast path vs shape-generic path) and launches.  No CPU fallback.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>

#include "slb_fast.cuh"
#include "slb_generic_env.cuh"
#include "slb_big_rows.cuh"
#include "slb_microbench.cuh"
#include "slb_emd.cuh"

using namespace slb;

#ifndef SLB_BIG_OPT_DEFAULT
#define SLB_BIG_OPT_DEFAULT 55
#endif

namespace {

thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launches{0};

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

#define SLB_CUDA(expr)                                                                       \
    do {                                                                                     \
        cudaError_t e_ = (expr);                                                             \
        if (e_ != cudaSuccess)                                                               \
            return fail(SLB_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_));       \
    } while (0)

#define SLB_LAUNCHED()                                                                       \
    do {                                                                                     \
        g_launches.fetch_add(1, std::memory_order_relaxed);                                  \
        cudaError_t e_ = cudaGetLastError();                                                 \
        if (e_ != cudaSuccess)                                                               \
            return fail(SLB_ERR_CUDA, "kernel launch failed: %s", cudaGetErrorString(e_));   \
    } while (0)

int check_shape(int n, int h, int w) {
    if (n < 0) return fail(SLB_ERR_INVALID, "n must be >= 0 (got %d)", n);
    if (h < 3 || w < 3)
        return fail(SLB_ERR_INVALID,
                    "board must be at least 3x3 (got %dx%d): the reference double-counts wrapped "
                    "neighbours below that (SURVEY.md Appendix A)", h, w);
    if ((long long)h * w > (1ll << 28)) return fail(SLB_ERR_UNSUPPORTED, "board too large");
    return 0;
}

int check_pool(const slb_pool *p) {
    if (!p) return fail(SLB_ERR_INVALID, "pool is NULL");
    if (p->n_levels <= 0) return fail(SLB_ERR_INVALID, "pool has no levels");
    if (p->words_per_row != (p->w + 31) / 32) return fail(SLB_ERR_INVALID, "pool.words_per_row");
    if (!p->board || !p->goals || !p->agent_x || !p->agent_y || !p->orientation ||
        !p->spawn_prob || !p->min_performance || !p->baseline || !p->exit_mask || !p->plane_bits ||
        !p->level_consts)
        return fail(SLB_ERR_INVALID, "pool has NULL arrays");
    if ((fast_shape(p->h, p->w) || big_shape(p->h, p->w)) && (!p->board_res || !p->goals_res))
        return fail(SLB_ERR_INVALID, "pool.board_res / goals_res are required for %dx%d boards", p->h, p->w);
    return check_shape(0, p->h, p->w);
}

constexpr size_t kSmemStatic = 2048;  // static shared memory of the generic kernels (upper bound)
size_t g_smem_optin = 0;

size_t smem_optin() {
    if (!g_smem_optin) {
        int dev = 0, v = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        g_smem_optin = (size_t)v;
    }
    return g_smem_optin;
}

template <typename K>
int allow_smem(K kernel, size_t bytes) {
    if (bytes > 48 * 1024)
        SLB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return 0;
}

// launches the big-board env kernel with the stages in a.ops: the register sliding-window variant where the
// shape has one (rows of 2 / 4 / 8 words, see slb_big_rows.cuh; SLB_BIG_ROWS=0 turns it off), else the
// shared-memory snapshot variant
int launch_big(const slb_pool *pool, BigArgs a, cudaStream_t st) {
    static const bool rows_on = []() { const char *v = getenv("SLB_BIG_ROWS"); return !v || v[0] != '0'; }();
    if (rows_on && rows_shape(pool->h, pool->w)) {
        const size_t bytes = rows_smem_bytes(pool->h, pool->w);
        const int wpr = pool->w / 32;
        // the fused step has its own kernel (slb_big_rows.cuh, rows_step_body); SLB_BIG_OPT picks one of the
        // compiled variants (0 = the all-stages kernel below), for A/B runs
        static const int opt = []() { const char *v = getenv("SLB_BIG_OPT"); return v ? atoi(v) : SLB_BIG_OPT_DEFAULT; }();
        if (a.ops == (BIG_ACTION | BIG_PHYSICS | BIG_BOOK) && opt != 0) {
#define SLB_ROWS_STEP(WPR_, OPT_)                                                              \
    do {                                                                                       \
        const size_t sbytes_ = rows_step_smem_bytes(pool->h, pool->w);                         \
        if (int rc = allow_smem(big_rows_step_kernel<WPR_, OPT_>, sbytes_)) return rc;         \
        cudaLaunchAttribute attr_[1];                                                          \
        attr_[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                      \
        attr_[0].val.programmaticStreamSerializationAllowed = 1;                               \
        cudaLaunchConfig_t cfg_ = {};                                                          \
        cfg_.gridDim = dim3(a.n); cfg_.blockDim = dim3(ROWS_THREADS);                          \
        cfg_.dynamicSmemBytes = sbytes_; cfg_.stream = st;                                     \
        cfg_.attrs = attr_; cfg_.numAttrs = ((OPT_) & RO_PDL) ? 1 : 0;                         \
        SLB_CUDA(cudaLaunchKernelEx(&cfg_, big_rows_step_kernel<WPR_, OPT_>, *pool, a));       \
        return SLB_OK;                                                                         \
    } while (0)
            if (wpr == 8) {
                switch (opt) {
                case 16: SLB_ROWS_STEP(8, 16);
                case 17: SLB_ROWS_STEP(8, 17);
                case 19: SLB_ROWS_STEP(8, 19);
                case 23: SLB_ROWS_STEP(8, 23);
                case 29: SLB_ROWS_STEP(8, 29);
                case 31: SLB_ROWS_STEP(8, 31);
                case 55: SLB_ROWS_STEP(8, 55);
                default: SLB_ROWS_STEP(8, 55);
                }
            } else if (wpr == 4) {
                SLB_ROWS_STEP(4, 55);
            } else {
                SLB_ROWS_STEP(2, 55);
            }
#undef SLB_ROWS_STEP
        }
        if (wpr == 8) {
            if (int rc = allow_smem(big_rows_env_kernel<8>, bytes)) return rc;
            big_rows_env_kernel<8><<<a.n, ROWS_THREADS, bytes, st>>>(*pool, a);
        } else if (wpr == 4) {
            if (int rc = allow_smem(big_rows_env_kernel<4>, bytes)) return rc;
            big_rows_env_kernel<4><<<a.n, ROWS_THREADS, bytes, st>>>(*pool, a);
        } else {
            if (int rc = allow_smem(big_rows_env_kernel<2>, bytes)) return rc;
            big_rows_env_kernel<2><<<a.n, ROWS_THREADS, bytes, st>>>(*pool, a);
        }
        return SLB_OK;
    }
    const size_t bytes = big_smem_bytes(pool->h, pool->w);
    if (int rc = allow_smem(big_env_kernel, bytes)) return rc;
    big_env_kernel<<<a.n, BIG_THREADS, bytes, st>>>(*pool, a);
    return SLB_OK;
}

BigArgs big_args(int n, void *board, void *goals, slb_env_state *state, int ops) {
    BigArgs a = {};
    a.board = (uint32_t *)board;
    a.goals = (uint32_t *)goals;
    a.state = state;
    a.n = n;
    a.ops = ops;
    return a;
}

}  // namespace

extern "C" {

const char *slb_last_error(void) { return g_err; }
int slb_abi_version(void) { return SLB_ABI_VERSION; }
uint64_t slb_launch_count(void) { return g_launches.load(); }

size_t slb_resident_bytes(int h, int w) {
    if (h < 3 || w < 3) return 0;
    if (fast_shape(h, w)) return resident_words(h) * 4;
    if (big_shape(h, w)) return (size_t)64 * h * (w / 32);
    return (size_t)h * w * 2;
}

static int refresh_groups(slb_pool *pool, cudaStream_t st) {
    uint32_t bits[2] = {0, 0};
    SLB_CUDA(cudaMemcpyAsync(bits, pool->plane_bits, sizeof(bits), cudaMemcpyDeviceToHost, st));
    SLB_CUDA(cudaStreamSynchronize(st));
    pool->plane_groups[0] = groups_of(bits[0]);
    pool->plane_groups[1] = groups_of(bits[1]);
    return SLB_OK;
}

int slb_resident_import(const uint16_t *d_cells, void *d_res, int n, int h, int w,
                        slb_pool *pool, int which, void *stream) {
    if (int rc = check_shape(n, h, w)) return rc;
    if (!d_cells || !d_res) return fail(SLB_ERR_INVALID, "slb_resident_import: NULL pointer");
    if (pool && (which < 0 || which > 1 || !pool->plane_bits || pool->h != h || pool->w != w))
        return fail(SLB_ERR_INVALID, "slb_resident_import: pool / which do not match the boards");
    uint32_t *pool_plane_bits = pool ? pool->plane_bits + which : nullptr;
    if (n == 0) return SLB_OK;
    // an env buffer written from outside: the fused step stops trusting its cached score / static goals
    if (pool && d_res != pool->board_res && d_res != pool->goals_res) pool->external_writes = 1;
    cudaStream_t st = (cudaStream_t)stream;
    if (big_shape(h, w)) {
        if ((uintptr_t)d_res % 16 || (uintptr_t)d_cells % 16)
            return fail(SLB_ERR_INVALID, "resident buffers and cell arrays must be 16-byte aligned");
        big_import_kernel<<<n, 256, 0, st>>>(d_cells, (uint32_t *)d_res, h, w, pool_plane_bits);
        SLB_LAUNCHED();
        if (pool) return refresh_groups(pool, st);
        return SLB_OK;
    }
    if (!fast_shape(h, w)) {
        SLB_CUDA(cudaMemcpyAsync(d_res, d_cells, (size_t)n * h * w * 2, cudaMemcpyDeviceToDevice, st));
        return SLB_OK;
    }
    if ((uintptr_t)d_res % 16) return fail(SLB_ERR_INVALID, "resident buffers must be 16-byte aligned");
    res_import_kernel<<<(n + RES_WARPS - 1) / RES_WARPS, RES_WARPS * 32, 0, st>>>(
        d_cells, (uint32_t *)d_res, n, h, w, pool_plane_bits, fast_vec_ok(h, w, d_cells, d_cells));
    SLB_LAUNCHED();
    if (pool) return refresh_groups(pool, st);
    return SLB_OK;
}

int slb_resident_export(const void *d_res, uint16_t *d_cells, int n, int h, int w, void *stream) {
    if (int rc = check_shape(n, h, w)) return rc;
    if (!d_cells || !d_res) return fail(SLB_ERR_INVALID, "slb_resident_export: NULL pointer");
    if (n == 0) return SLB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (big_shape(h, w)) {
        if ((uintptr_t)d_res % 16 || (uintptr_t)d_cells % 16)
            return fail(SLB_ERR_INVALID, "resident buffers and cell arrays must be 16-byte aligned");
        big_export_kernel<<<n, 256, 0, st>>>((const uint32_t *)d_res, d_cells, h, w);
        SLB_LAUNCHED();
        return SLB_OK;
    }
    if (!fast_shape(h, w)) {
        SLB_CUDA(cudaMemcpyAsync(d_cells, d_res, (size_t)n * h * w * 2, cudaMemcpyDeviceToDevice, st));
        return SLB_OK;
    }
    if ((uintptr_t)d_res % 16) return fail(SLB_ERR_INVALID, "resident buffers must be 16-byte aligned");
    res_export_kernel<<<(n + RES_WARPS - 1) / RES_WARPS, RES_WARPS * 32, 0, st>>>(
        (const uint32_t *)d_res, d_cells, n, h, w, fast_vec_ok(h, w, d_cells, d_cells));
    SLB_LAUNCHED();
    return SLB_OK;
}

int slb_pool_prepare(slb_pool *pool, void *stream) {
    if (int rc = check_pool(pool)) return rc;
    pool_prepare_kernel<<<pool->n_levels, GEN_THREADS, 0, (cudaStream_t)stream>>>(*pool);
    SLB_LAUNCHED();
    if (fast_shape(pool->h, pool->w) || big_shape(pool->h, pool->w)) {
        if (int rc = slb_resident_import(pool->board, pool->board_res, pool->n_levels, pool->h, pool->w,
                                         pool, 0, stream)) return rc;
        if (int rc = slb_resident_import(pool->goals, pool->goals_res, pool->n_levels, pool->h, pool->w,
                                         pool, 1, stream)) return rc;
    }
    if (fast_shape(pool->h, pool->w)) {
        res_level_flags_kernel<<<(pool->n_levels + RES_WARPS - 1) / RES_WARPS, RES_WARPS * 32, 0,
                                 (cudaStream_t)stream>>>(*pool);
        SLB_LAUNCHED();
    }
    pool->external_writes = 0;
    return SLB_OK;
}

static int advance_once(const uint16_t *in, uint16_t *out, int n, int h, int w, const float *p,
                        uint64_t seed, uint64_t env0, uint64_t *ctr, cudaStream_t st) {
    if (fast_shape(h, w)) {
        launch_advance_fast(in, out, n, h, w, p, seed, env0, ctr, 1, st);
        SLB_LAUNCHED();
        return SLB_OK;
    }
    size_t bytes = (size_t)h * w * 2;
    if (bytes + kSmemStatic <= smem_optin()) {
        if (int rc = allow_smem(advance_generic_kernel<true>, bytes)) return rc;
        advance_generic_kernel<true><<<n, GEN_THREADS, bytes, st>>>(in, out, h, w, p, seed, env0, ctr);
    } else {
        advance_generic_kernel<false><<<n, GEN_THREADS, 0, st>>>(in, out, h, w, p, seed, env0, ctr);
    }
    SLB_LAUNCHED();
    return SLB_OK;
}

int slb_advance_board(const uint16_t *d_in, uint16_t *d_out, int n, int h, int w,
                      const float *d_spawn_prob, uint64_t seed, uint64_t env_id0,
                      uint64_t *d_counters, int k_steps, uint16_t *d_scratch, void *stream) {
    if (int rc = check_shape(n, h, w)) return rc;
    if (!d_in || !d_out || !d_spawn_prob) return fail(SLB_ERR_INVALID, "NULL board/spawn_prob pointer");
    if (d_in == d_out) return fail(SLB_ERR_INVALID, "advance_board is out of place: d_in == d_out");
    if (k_steps < 1) return fail(SLB_ERR_INVALID, "k_steps must be >= 1");
    if (n == 0) return SLB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (fast_shape(h, w)) {  // k steps on chip, one HBM round trip
        launch_advance_fast(d_in, d_out, n, h, w, d_spawn_prob, seed, env_id0, d_counters, k_steps, st);
        SLB_LAUNCHED();
        return SLB_OK;
    }
    if (k_steps > 1 && !d_scratch)
        return fail(SLB_ERR_INVALID, "k_steps > 1 on a %dx%d board needs d_scratch", h, w);
    // ping-pong so that the last step lands in d_out
    const uint16_t *src = d_in;
    for (int k = 0; k < k_steps; k++) {
        uint16_t *dst = ((k_steps - 1 - k) % 2 == 0) ? d_out : d_scratch;
        if (int rc = advance_once(src, dst, n, h, w, d_spawn_prob, seed, env_id0, d_counters, st)) return rc;
        src = dst;
    }
    return SLB_OK;
}

// Per-thread staging of slb_advance_board_host: a growable device block, a growable pinned (and therefore
// device-mapped) host block and a non-blocking stream, all kept between calls — no allocation, no pageable
// copy and no legacy-default-stream synchronisation per call.
namespace {
struct AdvStage {
    int dev;
    cudaStream_t st;
    uint8_t *d, *h;
    size_t dcap, hcap;
};
thread_local AdvStage g_adv = {-1, nullptr, nullptr, nullptr, 0, 0};
constexpr size_t kAdvZeroCopyBytes = 64 * 1024;   // up to here the kernel reads / writes the pinned block itself
constexpr size_t kAdvPinnedBytes = 256u << 20;    // larger batches are copied straight from / to the caller's arrays

int adv_stage(size_t dbytes, size_t hbytes, AdvStage **out) {
    int dev = 0;
    SLB_CUDA(cudaGetDevice(&dev));
    AdvStage *a = &g_adv;
    if (a->dev != dev) {  // first use on this thread, or the thread moved to another device
        if (a->st) { cudaStreamSynchronize(a->st); cudaStreamDestroy(a->st); }
        if (a->d) cudaFree(a->d);
        if (a->h) cudaFreeHost(a->h);
        *a = AdvStage{dev, nullptr, nullptr, nullptr, 0, 0};
        SLB_CUDA(cudaStreamCreateWithFlags(&a->st, cudaStreamNonBlocking));
    }
    if (dbytes > a->dcap) {
        if (a->d) cudaFree(a->d);
        a->d = nullptr; a->dcap = 0;
        const size_t cap = dbytes + dbytes / 2;
        SLB_CUDA(cudaMalloc(&a->d, cap));
        a->dcap = cap;
    }
    if (hbytes > a->hcap) {
        if (a->h) cudaFreeHost(a->h);
        a->h = nullptr; a->hcap = 0;
        const size_t cap = hbytes + hbytes / 2;
        SLB_CUDA(cudaMallocHost(&a->h, cap));
        a->hcap = cap;
    }
    *out = a;
    return SLB_OK;
}
inline size_t up16(size_t x) { return (x + 15) & ~(size_t)15; }
}  // namespace

int slb_advance_board_host(const uint16_t *h_in, uint16_t *h_out, int n, int h, int w,
                           const float *h_spawn_prob, uint64_t seed, uint64_t env_id0,
                           uint64_t *h_counters, int k_steps) {
    if (int rc = check_shape(n, h, w)) return rc;
    if (!h_in || !h_out || !h_spawn_prob) return fail(SLB_ERR_INVALID, "NULL host pointer");
    if (k_steps < 1) return fail(SLB_ERR_INVALID, "k_steps must be >= 1");
    if (n == 0) return SLB_OK;
    const size_t cells = (size_t)n * h * w, bb = up16(cells * 2), pb = up16((size_t)n * 4), cb = up16((size_t)n * 8);
    const bool need_scratch = k_steps > 1 && !fast_shape(h, w);
    // block layout (host and device alike): [boards in | boards out | spawn_prob | counters | scratch]
    const size_t o_out = bb, o_p = 2 * bb, o_c = o_p + pb, o_s = o_c + cb, total = o_s + (need_scratch ? bb : 0);
    const bool zero_copy = 2 * bb <= kAdvZeroCopyBytes && !need_scratch;
    const bool pinned = o_s <= kAdvPinnedBytes;
    AdvStage *a = nullptr;
    if (int rc = adv_stage(zero_copy ? 16 : total, pinned ? o_s : 16, &a)) return rc;
    cudaStream_t st = a->st;
    int rc;
    if (zero_copy) {
        // a few boards: the kernel reads the inputs from and writes the results to the pinned block over
        // PCIe itself — one launch and one wait, no DMA copy
        memcpy(a->h, h_in, cells * 2);
        memcpy(a->h + o_p, h_spawn_prob, (size_t)n * 4);
        if (h_counters) memcpy(a->h + o_c, h_counters, (size_t)n * 8);
        else memset(a->h + o_c, 0, (size_t)n * 8);
        t_advance_no_bulk = true;
        rc = slb_advance_board((const uint16_t *)a->h, (uint16_t *)(a->h + o_out), n, h, w, (const float *)(a->h + o_p),
                               seed, env_id0, (uint64_t *)(a->h + o_c), k_steps, nullptr, st);
        t_advance_no_bulk = false;
        if (rc != SLB_OK) return rc;
        SLB_CUDA(cudaStreamSynchronize(st));
        memcpy(h_out, a->h + o_out, cells * 2);
        if (h_counters) memcpy(h_counters, a->h + o_c, (size_t)n * 8);
        return SLB_OK;
    }
    const uint8_t *src_in = (const uint8_t *)h_in, *src_p = (const uint8_t *)h_spawn_prob;
    if (pinned) {
        memcpy(a->h, h_in, cells * 2);
        memcpy(a->h + o_p, h_spawn_prob, (size_t)n * 4);
        src_in = a->h; src_p = a->h + o_p;
        if (h_counters) memcpy(a->h + o_c, h_counters, (size_t)n * 8);
    }
    SLB_CUDA(cudaMemcpyAsync(a->d, src_in, cells * 2, cudaMemcpyHostToDevice, st));
    SLB_CUDA(cudaMemcpyAsync(a->d + o_p, src_p, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    if (h_counters)
        SLB_CUDA(cudaMemcpyAsync(a->d + o_c, pinned ? (const void *)(a->h + o_c) : (const void *)h_counters,
                                 (size_t)n * 8, cudaMemcpyHostToDevice, st));
    else
        SLB_CUDA(cudaMemsetAsync(a->d + o_c, 0, (size_t)n * 8, st));
    rc = slb_advance_board((const uint16_t *)a->d, (uint16_t *)(a->d + o_out), n, h, w, (const float *)(a->d + o_p),
                           seed, env_id0, (uint64_t *)(a->d + o_c), k_steps,
                           need_scratch ? (uint16_t *)(a->d + o_s) : nullptr, st);
    if (rc != SLB_OK) return rc;
    SLB_CUDA(cudaMemcpyAsync(pinned ? (void *)(a->h + o_out) : (void *)h_out, a->d + o_out, cells * 2,
                             cudaMemcpyDeviceToHost, st));
    if (h_counters)
        SLB_CUDA(cudaMemcpyAsync(pinned ? (void *)(a->h + o_c) : (void *)h_counters, a->d + o_c, (size_t)n * 8,
                                 cudaMemcpyDeviceToHost, st));
    SLB_CUDA(cudaStreamSynchronize(st));
    if (pinned) {
        memcpy(h_out, a->h + o_out, cells * 2);
        if (h_counters) memcpy(h_counters, a->h + o_c, (size_t)n * 8);
    }
    return SLB_OK;
}

#define RES_GRID(n) (((n) + RES_WARPS - 1) / RES_WARPS), RES_WARPS * 32, 0

int slb_env_reset(const slb_pool *pool, int n, const int32_t *d_level_ids, const uint8_t *d_mask,
                  void *d_board, void *d_goals, slb_env_state *d_state, void *stream) {
    if (int rc = check_pool(pool)) return rc;
    if (n < 0 || !d_level_ids || !d_board || !d_goals || !d_state)
        return fail(SLB_ERR_INVALID, "slb_env_reset: bad arguments");
    if (n == 0) return SLB_OK;
    if (fast_shape(pool->h, pool->w)) {
        res_reset_kernel<<<RES_GRID(n), (cudaStream_t)stream>>>(*pool, n, d_level_ids, d_mask,
                                                              (uint32_t *)d_board, (uint32_t *)d_goals, d_state);
    } else if (big_shape(pool->h, pool->w)) {
        BigArgs a = big_args(n, d_board, d_goals, d_state, BIG_RESET);
        a.level_ids = d_level_ids;
        a.mask = d_mask;
        if (int rc = launch_big(pool, a, (cudaStream_t)stream)) return rc;
    } else
        reset_generic_kernel<<<n, GEN_THREADS, 0, (cudaStream_t)stream>>>(
            *pool, d_level_ids, d_mask, (uint16_t *)d_board, (uint16_t *)d_goals, d_state);
    SLB_LAUNCHED();
    return SLB_OK;
}

int slb_env_set_min_performance(int n, slb_env_state *d_state, const uint8_t *d_mask, double value,
                                void *stream) {
    if (n < 0 || !d_state) return fail(SLB_ERR_INVALID, "slb_env_set_min_performance: bad arguments");
    if (n == 0) return SLB_OK;
    set_min_performance_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(n, d_state, d_mask, value);
    SLB_LAUNCHED();
    return SLB_OK;
}

int slb_execute_action(slb_pool *pool, int n, void *d_board, slb_env_state *d_state,
                       const uint8_t *d_actions, int32_t *d_reward, void *stream) {
    if (int rc = check_pool(pool)) return rc;
    if (n < 0 || !d_board || !d_state || !d_actions)
        return fail(SLB_ERR_INVALID, "slb_execute_action: bad arguments");
    if (n == 0) return SLB_OK;
    pool->external_writes = 1;  // a board written outside the fused step (see slb_pool)
    if (fast_shape(pool->h, pool->w)) {
        res_action_kernel<<<RES_GRID(n), (cudaStream_t)stream>>>(*pool, n, (uint32_t *)d_board, d_state,
                                                               d_actions, d_reward);
    } else if (big_shape(pool->h, pool->w)) {
        BigArgs a = big_args(n, d_board, nullptr, d_state, BIG_ACTION);
        a.actions = d_actions;
        a.reward_out = d_reward;
        if (int rc = launch_big(pool, a, (cudaStream_t)stream)) return rc;
    } else
        execute_action_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
            (uint16_t *)d_board, n, pool->h, pool->w, d_state, d_actions, d_reward);
    SLB_LAUNCHED();
    return SLB_OK;
}

int slb_game_advance(slb_pool *pool, int n, void *d_board, void *d_goals,
                     slb_env_state *d_state, uint64_t seed, uint64_t env_id0, void *stream) {
    if (int rc = check_pool(pool)) return rc;
    if (n < 0 || !d_board || !d_goals || !d_state)
        return fail(SLB_ERR_INVALID, "slb_game_advance: bad arguments");
    if (n == 0) return SLB_OK;
    pool->external_writes = 1;  // a board written outside the fused step (see slb_pool)
    cudaStream_t st = (cudaStream_t)stream;
    if (fast_shape(pool->h, pool->w)) {
        res_advance_kernel<<<RES_GRID(n), st>>>(*pool, n, (uint32_t *)d_board, (uint32_t *)d_goals,
                                               d_state, seed, env_id0);
    } else if (big_shape(pool->h, pool->w)) {
        BigArgs a = big_args(n, d_board, d_goals, d_state, BIG_PHYSICS);
        a.seed = seed;
        a.env_id0 = env_id0;
        if (int rc = launch_big(pool, a, st)) return rc;
    } else {
        size_t bytes = (size_t)pool->h * pool->w * 2;
        if (bytes + kSmemStatic > smem_optin())
            return fail(SLB_ERR_UNSUPPORTED, "game advance needs the %dx%d board in shared memory",
                        pool->h, pool->w);
        if (int rc = allow_smem(game_advance_generic_kernel, bytes)) return rc;
        game_advance_generic_kernel<<<n, GEN_THREADS, bytes, st>>>(
            *pool, (uint16_t *)d_board, (uint16_t *)d_goals, d_state, seed, env_id0);
    }
    SLB_LAUNCHED();
    return SLB_OK;
}

int slb_points_and_performance(const slb_pool *pool, int n, const void *d_board,
                               const void *d_goals, const slb_env_state *d_state,
                               int32_t *d_points, int32_t *d_completed, int32_t *d_possible,
                               uint8_t *d_can_exit, void *stream) {
    if (int rc = check_pool(pool)) return rc;
    if (n < 0 || !d_board || !d_goals || !d_state) return fail(SLB_ERR_INVALID, "slb_points: bad arguments");
    if (n == 0) return SLB_OK;
    if (fast_shape(pool->h, pool->w))
        res_score_kernel<false><<<RES_GRID(n), (cudaStream_t)stream>>>(
            *pool, n, (uint32_t *)d_board, (const uint32_t *)d_goals, (slb_env_state *)d_state, d_points,
            d_completed, d_possible, d_can_exit);
    else if (big_shape(pool->h, pool->w)) {
        BigArgs a = big_args(n, (void *)d_board, (void *)d_goals, (slb_env_state *)d_state, BIG_POINTS);
        a.points = d_points; a.completed = d_completed; a.possible = d_possible; a.can_exit = d_can_exit;
        if (int rc = launch_big(pool, a, (cudaStream_t)stream)) return rc;
    } else
        points_generic_kernel<<<n, GEN_THREADS, 0, (cudaStream_t)stream>>>(
            *pool, (const uint16_t *)d_board, (const uint16_t *)d_goals, d_state, d_points, d_completed,
            d_possible, d_can_exit);
    SLB_LAUNCHED();
    return SLB_OK;
}

int slb_update_exit_colors(slb_pool *pool, int n, void *d_board, const void *d_goals,
                           slb_env_state *d_state, void *stream) {
    if (int rc = check_pool(pool)) return rc;
    if (n < 0 || !d_board || !d_goals || !d_state)
        return fail(SLB_ERR_INVALID, "slb_update_exit_colors: bad arguments");
    if (n == 0) return SLB_OK;
    pool->external_writes = 1;  // a board written outside the fused step (see slb_pool)
    if (fast_shape(pool->h, pool->w))
        res_score_kernel<true><<<RES_GRID(n), (cudaStream_t)stream>>>(
            *pool, n, (uint32_t *)d_board, (const uint32_t *)d_goals, d_state, nullptr, nullptr, nullptr, nullptr);
    else if (big_shape(pool->h, pool->w)) {
        BigArgs a = big_args(n, d_board, (void *)d_goals, d_state, BIG_PAINT);
        if (int rc = launch_big(pool, a, (cudaStream_t)stream)) return rc;
    } else
        exit_colors_generic_kernel<<<n, GEN_THREADS, 0, (cudaStream_t)stream>>>(
            *pool, (uint16_t *)d_board, (const uint16_t *)d_goals, d_state);
    SLB_LAUNCHED();
    return SLB_OK;
}

static int env_step_impl(const slb_pool *pool, int n, void *d_board, void *d_goals,
                         slb_env_state *d_state, const uint8_t *d_actions, int time_limit, uint64_t seed,
                         uint64_t env_id0, int32_t *d_reward, uint8_t *d_done, int auto_reset_stride,
                         void *stream, unsigned long long *gc) {
    if (int rc = check_pool(pool)) return rc;
    if (n < 0 || !d_board || !d_goals || !d_state || !d_actions)
        return fail(SLB_ERR_INVALID, "slb_env_step: bad arguments");
    if (n == 0) return SLB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (fast_shape(pool->h, pool->w)) {
        if ((uint64_t)n * 16u * (uint64_t)pool->h >= (1ull << 32))
            return fail(SLB_ERR_UNSUPPORTED, "too many envs per launch (resident word offsets are 32-bit)");
        launch_res_env_rollout(*pool, n, (uint32_t *)d_board, (uint32_t *)d_goals, d_state, d_actions, 1,
                               time_limit, seed, env_id0, d_reward, d_done, auto_reset_stride, st, gc);
        SLB_LAUNCHED();
        return SLB_OK;
    }
    if (gc) return fail(SLB_ERR_UNSUPPORTED, "the in-kernel global counter needs boards of 4..32 rows / columns");
    if (big_shape(pool->h, pool->w)) {
        BigArgs a = big_args(n, d_board, d_goals, d_state, BIG_ACTION | BIG_PHYSICS | BIG_BOOK);
        a.actions = d_actions; a.time_limit = time_limit; a.seed = seed; a.env_id0 = env_id0;
        a.reward_out = d_reward; a.done_out = d_done; a.auto_reset_stride = auto_reset_stride;
        if (int rc = launch_big(pool, a, st)) return rc;
        SLB_LAUNCHED();
        return SLB_OK;
    }
    size_t bytes = (size_t)pool->h * pool->w * 2;
    if (bytes + kSmemStatic > smem_optin())
        return fail(SLB_ERR_UNSUPPORTED, "fused env step needs the %dx%d board in shared memory",
                    pool->h, pool->w);
    if (int rc = allow_smem(env_step_generic_kernel, bytes)) return rc;
    env_step_generic_kernel<<<n, GEN_THREADS, bytes, st>>>(*pool, (uint16_t *)d_board, (uint16_t *)d_goals, d_state, d_actions,
                                                          time_limit, seed, env_id0, d_reward,
                                                          d_done, auto_reset_stride);
    SLB_LAUNCHED();
    return SLB_OK;
}

int slb_env_step(const slb_pool *pool, int n, void *d_board, void *d_goals,
                 slb_env_state *d_state, const uint8_t *d_actions, int time_limit, uint64_t seed,
                 uint64_t env_id0, int32_t *d_reward, uint8_t *d_done, int auto_reset_stride,
                 void *stream) {
    return env_step_impl(pool, n, d_board, d_goals, d_state, d_actions, time_limit, seed, env_id0, d_reward,
                         d_done, auto_reset_stride, stream, nullptr);
}

int slb_env_rollout(const slb_pool *pool, int n, void *d_board, void *d_goals,
                    slb_env_state *d_state, const uint8_t *d_actions, int t_steps, int time_limit,
                    uint64_t seed, uint64_t env_id0, int32_t *d_reward, uint8_t *d_done,
                    int auto_reset_stride, void *stream) {
    if (int rc = check_pool(pool)) return rc;
    if (n < 0 || t_steps < 0 || !d_board || !d_goals || !d_state || !d_actions)
        return fail(SLB_ERR_INVALID, "slb_env_rollout: bad arguments");
    if (n == 0 || t_steps == 0) return SLB_OK;
    if (fast_shape(pool->h, pool->w)) {
        if ((uint64_t)n * 16u * (uint64_t)pool->h >= (1ull << 32))
            return fail(SLB_ERR_UNSUPPORTED, "too many envs per launch (resident word offsets are 32-bit)");
        launch_res_env_rollout(*pool, n, (uint32_t *)d_board, (uint32_t *)d_goals, d_state, d_actions,
                               t_steps, time_limit, seed, env_id0, d_reward, d_done, auto_reset_stride,
                               (cudaStream_t)stream);
        SLB_LAUNCHED();
        return SLB_OK;
    }
    for (int t = 0; t < t_steps; t++) {
        int rc = slb_env_step(pool, n, d_board, d_goals, d_state, d_actions + (size_t)t * n, time_limit,
                              seed, env_id0, d_reward ? d_reward + (size_t)t * n : nullptr,
                              d_done ? d_done + (size_t)t * n : nullptr, auto_reset_stride, stream);
        if (rc) return rc;
    }
    return SLB_OK;
}


// ---- host-buffer steps ---------------------------------------------------------------------------
// One slot per (calling thread, stream): a device staging block [reward 4n | done n | pad | actions n]
// and a pinned host mirror of it.  A step runs as ONE instantiated CUDA graph {H2D actions, fused step,
// D2H reward+done} between those blocks, so that it costs the host one cudaGraphLaunch instead of two
// copies and a kernel launch; the caller's buffers are filled from the pinned mirror by
// slb_stream_wait (they need not be page-locked).  A slb_stepper owns its graph; the free function
// slb_env_step_host_async keeps a small least-recently-used cache of graphs per slot, keyed by the
// whole argument set.  SLB_HOST_GRAPH=0, the legacy default stream or a stream that is being captured
// take the plain path: copies straight from / to the caller's buffers.
namespace {
struct StepKey {
    slb_pool pool;
    int n, time_limit, stride, pad;
    void *board, *goals, *state;
    uint64_t seed, env_id0;
};
struct StepGraph {
    StepKey key;
    cudaGraphExec_t exec;
    uint64_t stamp;  // last use (LRU)
    uint8_t *d, *h;  // the staging blocks the graph copies between (a slot's, see below)
};
constexpr int kGraphsPerSlot = 32;
constexpr size_t kIoHeader = 16;  // in front of both blocks: device ticket counter / host sequence word
struct HostIo {
    cudaStream_t st;
    uint8_t *d, *h;  // the data parts (allocation + kIoHeader)
    size_t cap;
    StepGraph graphs[kGraphsPerSlot];
    int n_graphs;
    uint64_t clock;
    uint64_t last_use;
    int32_t *pend_reward;  // outputs of the step in flight still to be copied out of `h`
    uint8_t *pend_done;
    int pend_n;
    volatile uint32_t *pend_flag;  // direct path: the host word the step's last CTA publishes `pend_seq` to
    uint32_t pend_seq, seq;
};
constexpr int kMaxSlots = 16;  // streams with a host step in flight per calling thread
thread_local HostIo g_slots[kMaxSlots] = {};
thread_local int g_slots_used = 0;

int finish_pending(HostIo *s);

HostIo *find_slot(cudaStream_t st) {
    for (int k = 0; k < g_slots_used; k++)
        if (g_slots[k].st == st) return &g_slots[k];
    return nullptr;
}

void release_slot(HostIo *s) {
    for (int k = 0; k < s->n_graphs; k++) cudaGraphExecDestroy(s->graphs[k].exec);
    if (s->d) cudaFree(s->d - kIoHeader);
    if (s->h) cudaFreeHost(s->h - kIoHeader);
    memset(s, 0, sizeof(*s));
}

int host_io_slot(cudaStream_t st, size_t need, HostIo **out) {
    HostIo *s = find_slot(st);
    if (!s) {
        if (g_slots_used < kMaxSlots) {
            s = &g_slots[g_slots_used++];
        } else {
            // all slots taken: recycle the one used longest ago that has no step in flight (streams come
            // and go in a long-running process; a destroyed stream's slot is idle by construction)
            for (int k = 0; k < kMaxSlots; k++)
                if (!g_slots[k].pend_n && (!s || g_slots[k].last_use < s->last_use)) s = &g_slots[k];
            if (!s) return fail(SLB_ERR_UNSUPPORTED, "slb_env_step_host: more than 16 streams with a step in flight on one thread");
            release_slot(s);
        }
        memset(s, 0, sizeof(*s));
        s->st = st;
    }
    static thread_local uint64_t use_clock = 0;
    s->last_use = ++use_clock;
    if (need > s->cap) {
        // a step still in flight on this slot refers to the old blocks: deliver its outputs first
        if (s->pend_n)
            if (int rc = finish_pending(s)) return rc;
        SLB_CUDA(cudaStreamSynchronize(st));
        for (int k = 0; k < s->n_graphs; k++) cudaGraphExecDestroy(s->graphs[k].exec);
        s->n_graphs = 0;  // (graphs owned by steppers notice the new blocks and rebuild)
        if (s->d) cudaFree(s->d - kIoHeader);
        if (s->h) cudaFreeHost(s->h - kIoHeader);
        s->d = s->h = nullptr; s->cap = 0;
        uint8_t *db = nullptr, *hb = nullptr;
        SLB_CUDA(cudaMalloc(&db, need + kIoHeader));
        if (cudaMallocHost(&hb, need + kIoHeader) != cudaSuccess) {
            cudaFree(db);
            return fail(SLB_ERR_CUDA, "cudaMallocHost failed for the host-step staging block");
        }
        // the direct path keeps a ticket counter (device) and the last published step (host) in the headers
        if (cudaMemset(db, 0, need + kIoHeader) != cudaSuccess) {  // synchronous, not tied to `st`
            cudaFree(db); cudaFreeHost(hb);
            return fail(SLB_ERR_CUDA, "cudaMemset failed for the host-step staging block");
        }
        memset(hb, 0, need + kIoHeader);
        s->d = db + kIoHeader; s->h = hb + kIoHeader;
        s->cap = need;
    }
    *out = s;
    return SLB_OK;
}

int finish_pending(HostIo *s) {
    cudaError_t q;  // spin instead of a blocking wait: a step takes a few microseconds
    if (s->pend_flag) {
        // direct path: the step's last CTA wrote the outputs to `h` and then its sequence number; look at
        // the stream only now and then (a failed launch would never publish)
        for (unsigned spins = 0; *s->pend_flag != s->pend_seq; spins++) {
            if ((spins & 0xfff) == 0xfff) {
                q = cudaStreamQuery(s->st);
                if (q != cudaSuccess && q != cudaErrorNotReady) {
                    s->pend_flag = nullptr; s->pend_n = 0;
                    return fail(SLB_ERR_CUDA, "stream wait: %s", cudaGetErrorString(q));
                }
            }
        }
        __atomic_thread_fence(__ATOMIC_ACQUIRE);
        s->pend_flag = nullptr;
    } else {
        while ((q = cudaStreamQuery(s->st)) == cudaErrorNotReady) {}
        if (q != cudaSuccess) return fail(SLB_ERR_CUDA, "stream wait: %s", cudaGetErrorString(q));
    }
    if (s->pend_n) {
        if (s->pend_reward) memcpy(s->pend_reward, s->h, (size_t)4 * s->pend_n);
        if (s->pend_done) memcpy(s->pend_done, s->h + (size_t)4 * s->pend_n, (size_t)s->pend_n);
        s->pend_n = 0;
    }
    return SLB_OK;
}

bool host_graphs_enabled() {
    static const bool on = []() { const char *v = getenv("SLB_HOST_GRAPH"); return !v || v[0] != '0'; }();
    return on;
}

inline size_t io_off_act(int n) { return ((size_t)5 * n + 15) & ~(size_t)15; }

bool host_direct_enabled() {
    static const bool on = []() { const char *v = getenv("SLB_HOST_DIRECT"); return !v || v[0] != '0'; }();
    return on;
}

// capture {H2D actions, step, D2H reward+done} on the slot's stream and instantiate it
int build_step_graph(HostIo *s, const StepKey &k, cudaGraphExec_t *exec) {
    const size_t off_done = (size_t)4 * k.n, off_act = io_off_act(k.n);
    cudaGraph_t graph = nullptr;
    SLB_CUDA(cudaStreamBeginCapture(s->st, cudaStreamCaptureModeThreadLocal));
    cudaError_t e1 = cudaMemcpyAsync(s->d + off_act, s->h + off_act, (size_t)k.n, cudaMemcpyHostToDevice, s->st);
    int rc = slb_env_step(&k.pool, k.n, k.board, k.goals, (slb_env_state *)k.state, s->d + off_act, k.time_limit,
                          k.seed, k.env_id0, (int32_t *)s->d, s->d + off_done, k.stride, (void *)s->st);
    cudaError_t e2 = cudaMemcpyAsync(s->h, s->d, (size_t)5 * k.n, cudaMemcpyDeviceToHost, s->st);
    cudaError_t e3 = cudaStreamEndCapture(s->st, &graph);
    if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
    g_launches.fetch_sub(1, std::memory_order_relaxed);  // captured, not launched
    if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess || !graph) {
        if (graph) cudaGraphDestroy(graph);
        cudaGetLastError();
        return fail(SLB_ERR_CUDA, "slb_env_step_host: graph capture failed");
    }
    cudaError_t e4 = cudaGraphInstantiate(exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e4 != cudaSuccess) return fail(SLB_ERR_CUDA, "cudaGraphInstantiate failed: %s", cudaGetErrorString(e4));
    static const bool dbg = getenv("SLB_DEBUG") != nullptr;
    if (dbg) fprintf(stderr, "[slb] step graph built: stream %p, n=%d\n", (void *)s->st, k.n);
    return SLB_OK;
}

// `own`: the caller's graph (a stepper's), or NULL to use the slot's cache
int host_step(const slb_pool *pool, int n, void *d_board, void *d_goals, slb_env_state *d_state,
              const uint8_t *h_actions, int time_limit, uint64_t seed, uint64_t env_id0, int32_t *h_reward,
              uint8_t *h_done, int auto_reset_stride, void *stream, StepGraph *own, uint8_t *bound_dev = nullptr) {
    if (n < 0 || !h_actions) return fail(SLB_ERR_INVALID, "slb_env_step_host: bad arguments");
    if (n == 0) return SLB_OK;
    if (int rc = check_pool(pool)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t off_done = (size_t)4 * n, off_act = io_off_act(n);
    HostIo *s = nullptr;
    if (int rc = host_io_slot(st, off_act + (size_t)n, &s)) return rc;
    if (s->pend_n)  // an earlier step on this stream was never waited for: deliver its outputs first
        if (int rc = finish_pending(s)) return rc;
    uint8_t *d_io = s->d;

    bool graphed = host_graphs_enabled() && st != nullptr && st != cudaStreamLegacy && st != cudaStreamPerThread;
    if (graphed) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        SLB_CUDA(cudaStreamIsCapturing(st, &cs));
        graphed = cs == cudaStreamCaptureStatusNone;
    }
    if (graphed && host_direct_enabled() && n <= HOST_STEP_MAX_ENVS && fast_shape(pool->h, pool->w)) {
        // no DMA copy on either side: actions by value in the launch parameters, outputs written to the
        // pinned mirror by the step's last CTA, which then publishes this step's sequence number
        static thread_local ActBlock acts;
        memcpy(acts.a, h_actions, (size_t)n);
        HostTail ht;
        ht.d_reward = (int32_t *)d_io; ht.d_done = d_io + off_done;
        ht.counter = (unsigned *)(s->d - kIoHeader);
        ht.h_block = s->h; ht.h_seq = (uint32_t *)(s->h - kIoHeader);
        // outputs bound with slb_stepper_bind_outputs: the last CTA writes straight into the caller's page-locked
        // block [reward 4n | done n] and the copy out of the slot's mirror is skipped
        const bool zero_copy = bound_dev != nullptr;
        if (zero_copy) ht.h_block = bound_dev;
        ht.seq = ++s->seq;
        if (ht.seq == 0) ht.seq = ++s->seq;  // 0 is the freshly allocated block's value
        launch_res_env_step_host(*pool, n, (uint32_t *)d_board, (uint32_t *)d_goals, d_state, acts, time_limit,
                                 seed, env_id0, auto_reset_stride, ht, st);
        SLB_LAUNCHED();
        s->pend_reward = zero_copy ? nullptr : h_reward; s->pend_done = zero_copy ? nullptr : h_done; s->pend_n = n;
        s->pend_flag = (volatile uint32_t *)ht.h_seq; s->pend_seq = ht.seq;
        return SLB_OK;
    }
    if (!graphed) {
        SLB_CUDA(cudaMemcpyAsync(d_io + off_act, h_actions, (size_t)n, cudaMemcpyHostToDevice, st));
        int rc = slb_env_step(pool, n, d_board, d_goals, d_state, d_io + off_act, time_limit, seed, env_id0,
                              (int32_t *)d_io, d_io + off_done, auto_reset_stride, stream);
        if (rc) return rc;
        if (h_reward && h_done == (uint8_t *)(h_reward + n)) {  // adjacent host outputs: one copy
            SLB_CUDA(cudaMemcpyAsync(h_reward, d_io, (size_t)5 * n, cudaMemcpyDeviceToHost, st));
        } else {
            if (h_done) SLB_CUDA(cudaMemcpyAsync(h_done, d_io + off_done, (size_t)n, cudaMemcpyDeviceToHost, st));
            if (h_reward) SLB_CUDA(cudaMemcpyAsync(h_reward, d_io, (size_t)4 * n, cudaMemcpyDeviceToHost, st));
        }
        return SLB_OK;
    }

    StepKey key;
    memset(&key, 0, sizeof(key));
    key.pool = *pool; key.n = n; key.time_limit = time_limit; key.stride = auto_reset_stride;
    key.board = d_board; key.goals = d_goals; key.state = d_state; key.seed = seed; key.env_id0 = env_id0;
    StepGraph *hit = nullptr;
    if (own) {
        if (own->exec && own->d == s->d && own->h == s->h && !memcmp(&own->key, &key, sizeof(key))) {
            hit = own;
        } else {
            if (own->exec) { cudaGraphExecDestroy(own->exec); own->exec = nullptr; }
            if (int rc = build_step_graph(s, key, &own->exec)) { own->exec = nullptr; return rc; }
            own->key = key; own->d = s->d; own->h = s->h;
            hit = own;
        }
    } else {
        for (int k = 0; k < s->n_graphs; k++)
            if (!memcmp(&s->graphs[k].key, &key, sizeof(key))) { hit = &s->graphs[k]; break; }
        if (!hit) {
            cudaGraphExec_t exec = nullptr;
            if (int rc = build_step_graph(s, key, &exec)) return rc;
            if (s->n_graphs < kGraphsPerSlot) {
                hit = &s->graphs[s->n_graphs++];
            } else {  // replace the least recently used one
                hit = &s->graphs[0];
                for (int k = 1; k < kGraphsPerSlot; k++)
                    if (s->graphs[k].stamp < hit->stamp) hit = &s->graphs[k];
                cudaGraphExecDestroy(hit->exec);
            }
            hit->exec = exec; hit->key = key; hit->d = s->d; hit->h = s->h;
        }
        hit->stamp = ++s->clock;
    }
    memcpy(s->h + off_act, h_actions, (size_t)n);
    SLB_CUDA(cudaGraphLaunch(hit->exec, st));
    g_launches.fetch_add(1, std::memory_order_relaxed);  // the step kernel inside the graph
    s->pend_reward = h_reward; s->pend_done = h_done; s->pend_n = n;
    return SLB_OK;
}
}  // namespace

int slb_env_step_host_async(const slb_pool *pool, int n, void *d_board, void *d_goals,
                            slb_env_state *d_state, const uint8_t *h_actions, int time_limit,
                            uint64_t seed, uint64_t env_id0, int32_t *h_reward, uint8_t *h_done,
                            int auto_reset_stride, void *stream) {
    return host_step(pool, n, d_board, d_goals, d_state, h_actions, time_limit, seed, env_id0, h_reward,
                     h_done, auto_reset_stride, stream, nullptr);
}

int slb_env_step_host(const slb_pool *pool, int n, void *d_board, void *d_goals,
                      slb_env_state *d_state, const uint8_t *h_actions, int time_limit,
                      uint64_t seed, uint64_t env_id0, int32_t *h_reward, uint8_t *h_done,
                      int auto_reset_stride, void *stream) {
    if (int rc = slb_env_step_host_async(pool, n, d_board, d_goals, d_state, h_actions, time_limit, seed,
                                         env_id0, h_reward, h_done, auto_reset_stride, stream))
        return rc;
    return slb_stream_wait(stream);
}

int slb_stream_wait(void *stream) {
    if (HostIo *s = find_slot((cudaStream_t)stream)) return finish_pending(s);
    cudaError_t q;
    while ((q = cudaStreamQuery((cudaStream_t)stream)) == cudaErrorNotReady) {}
    if (q != cudaSuccess) return fail(SLB_ERR_CUDA, "stream wait: %s", cudaGetErrorString(q));
    return SLB_OK;
}

struct slb_stepper {
    const slb_pool *pool;
    int n, time_limit, stride;
    void *board, *goals;
    slb_env_state *state;
    uint64_t seed, env_id0;
    void *stream;
    int device;  // current device at creation; made current again by every step (callers hop devices)
    StepGraph graph;
    int32_t *bound_reward;  // slb_stepper_bind_outputs: the caller's page-locked output block and its device address
    uint8_t *bound_dev;
};

int slb_stepper_create(const slb_pool *pool, int n, void *d_board, void *d_goals, slb_env_state *d_state,
                       int time_limit, uint64_t seed, uint64_t env_id0, int auto_reset_stride,
                       void *stream, slb_stepper **out) {
    if (int rc = check_pool(pool)) return rc;
    if (!out || n < 0 || !d_board || !d_goals || !d_state)
        return fail(SLB_ERR_INVALID, "slb_stepper_create: bad arguments");
    int dev = 0;
    SLB_CUDA(cudaGetDevice(&dev));
    slb_stepper *s = new slb_stepper{pool, n, time_limit, auto_reset_stride, d_board, d_goals, d_state,
                                     seed, env_id0, stream, dev, {}, nullptr, nullptr};
    memset(&s->graph, 0, sizeof(s->graph));
    *out = s;
    return SLB_OK;
}

int slb_stepper_step(slb_stepper *s, const uint8_t *h_actions, int32_t *h_reward, uint8_t *h_done) {
    if (!s) return fail(SLB_ERR_INVALID, "stepper is NULL");
    int dev = s->device;
    if (cudaGetDevice(&dev) == cudaSuccess && dev != s->device) SLB_CUDA(cudaSetDevice(s->device));
    uint8_t *zc = (s->bound_dev && h_reward == s->bound_reward && h_done == (uint8_t *)(h_reward + s->n)) ? s->bound_dev
                                                                                                       : nullptr;
    return host_step(s->pool, s->n, s->board, s->goals, s->state, h_actions, s->time_limit, s->seed,
                     s->env_id0, h_reward, h_done, s->stride, s->stream, &s->graph, zc);
}

int slb_stepper_bind_outputs(slb_stepper *s, int32_t *h_reward, uint8_t *h_done) {
    if (!s) return fail(SLB_ERR_INVALID, "stepper is NULL");
    s->bound_reward = nullptr; s->bound_dev = nullptr;
    if (!h_reward && !h_done) return SLB_OK;  // unbind
    const size_t bytes = (size_t)5 * s->n;
    if (!h_reward || h_done != (uint8_t *)(h_reward + s->n) || (s->n & 15) || ((uintptr_t)h_reward & 15))
        return fail(SLB_ERR_INVALID, "slb_stepper_bind_outputs: the outputs must be ONE 16-byte aligned block "
                    "[int32 reward x n | uint8 done x n] with n a multiple of 16");
    cudaPointerAttributes a0, a1;
    if (cudaPointerGetAttributes(&a0, h_reward) != cudaSuccess ||
        cudaPointerGetAttributes(&a1, (const char *)h_reward + bytes - 1) != cudaSuccess ||
        a0.type != cudaMemoryTypeHost || a1.type != cudaMemoryTypeHost || !a0.devicePointer) {
        cudaGetLastError();
        return fail(SLB_ERR_INVALID, "slb_stepper_bind_outputs: the block is not page-locked host memory "
                    "(cudaHostAlloc / cudaHostRegister / torch pin_memory)");
    }
    s->bound_reward = h_reward;
    s->bound_dev = (uint8_t *)a0.devicePointer;
    return SLB_OK;
}

int slb_stepper_wait(slb_stepper *s) {
    if (!s) return fail(SLB_ERR_INVALID, "stepper is NULL");
    return slb_stream_wait(s->stream);
}

void slb_stepper_destroy(slb_stepper *s) {
    if (!s) return;
    if (s->graph.exec) cudaGraphExecDestroy(s->graph.exec);
    delete s;
}

int slb_side_effect_counts(const uint16_t *d_board0, const uint16_t *d_board1, int n, int h, int w,
                           const int32_t *d_num_steps, int num_samples, const float *d_spawn_prob,
                           uint64_t seed, uint64_t env_id0, uint64_t *d_counters, int kmax,
                           uint16_t *d_keys, uint16_t *d_counts, int32_t *d_nkeys,
                           uint16_t *d_scratch, void *stream) {
    if (int rc = check_shape(n, h, w)) return rc;
    if (!d_board0 || !d_board1 || !d_num_steps || !d_spawn_prob || !d_keys || !d_counts || !d_nkeys)
        return fail(SLB_ERR_INVALID, "slb_side_effect_counts: NULL pointer");
    if (kmax < 1 || kmax > SE_MAX_KEYS)
        return fail(SLB_ERR_INVALID, "kmax must be in [1, %d]", SE_MAX_KEYS);
    if (num_samples < 0 || num_samples > 65535)
        return fail(SLB_ERR_INVALID, "num_samples must fit the uint16 count maps (<= 65535)");
    if (n == 0) return SLB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (fast_shape(h, w) &&
        launch_side_effect_fast(d_board0, d_board1, n, h, w, d_num_steps, num_samples, d_spawn_prob,
                                seed, env_id0, d_counters, kmax, d_keys, d_counts, d_nkeys,
                                smem_optin() - kSmemStatic, st)) {
        SLB_LAUNCHED();
        return SLB_OK;
    }
    size_t bytes = (size_t)4 * h * w * 2;
    if (bytes + kSmemStatic <= smem_optin()) {
        if (int rc = allow_smem(side_effect_generic_kernel<true>, bytes)) return rc;
        side_effect_generic_kernel<true><<<n, GEN_THREADS, bytes, st>>>(
            d_board0, d_board1, h, w, d_num_steps, num_samples, d_spawn_prob, seed, env_id0,
            d_counters, kmax, d_keys, d_counts, d_nkeys, nullptr);
    } else {
        if (!d_scratch)
            return fail(SLB_ERR_INVALID, "%dx%d boards need d_scratch [N,4,H,W] for side effects", h, w);
        side_effect_generic_kernel<false><<<n, GEN_THREADS, 0, st>>>(
            d_board0, d_board1, h, w, d_num_steps, num_samples, d_spawn_prob, seed, env_id0,
            d_counters, kmax, d_keys, d_counts, d_nkeys, d_scratch);
    }
    SLB_LAUNCHED();
    return SLB_OK;
}

// ---- the EMD stage of side_effect_score, batched (side_effects.py:12-56 -> pyemd.emd) --------------
static int emd_args(const uint16_t *d_counts, const uint16_t *d_keys, int n, int kmax, int h, int w,
                    int num_samples, EmdArgs *a) {
    if (int rc = check_shape(n, h, w)) return rc;
    if (!d_counts || !d_keys || kmax < 1 || num_samples < 1)
        return fail(SLB_ERR_INVALID, "slb_emd: bad arguments");
    if (h > 256 || w > 256) return fail(SLB_ERR_UNSUPPORTED, "slb_emd: boards up to 256 x 256");
    if ((long long)n * kmax > (1ll << 30)) return fail(SLB_ERR_UNSUPPORTED, "slb_emd: too many problems");
    *a = EmdArgs{};
    a->counts = d_counts; a->keys = d_keys; a->n = n; a->kmax = kmax; a->h = h; a->w = w;
    a->num_samples = (double)num_samples;
    return SLB_OK;
}

int slb_emd_sizes(const uint16_t *d_counts, const uint16_t *d_keys, int n, int kmax, int h, int w,
                  int num_samples, int32_t *d_n_src, int32_t *d_n_snk, void *stream) {
    EmdArgs a;
    if (int rc = emd_args(d_counts, d_keys, n, kmax, h, w, num_samples, &a)) return rc;
    if (!d_n_src || !d_n_snk) return fail(SLB_ERR_INVALID, "slb_emd_sizes: NULL output");
    if (n == 0) return SLB_OK;
    const int probs = n * kmax;
    emd_sizes_kernel<<<(probs + EMD_WARPS - 1) / EMD_WARPS, EMD_WARPS * 32, 0, (cudaStream_t)stream>>>(a, d_n_src, d_n_snk);
    SLB_LAUNCHED();
    return SLB_OK;
}

int slb_emd_solve(const uint16_t *d_counts, const uint16_t *d_keys, int n, int kmax, int h, int w,
                  int num_samples, const double *d_lut, int wrap_x, int wrap_y, double extra_mass_penalty,
                  const int32_t *d_order, int n_problems, int vmax, size_t ft_stride, double *d_scratch,
                  size_t scratch_doubles, uint32_t *d_queue, double *d_emd, int32_t *d_status, void *stream) {
    EmdArgs a;
    if (int rc = emd_args(d_counts, d_keys, n, kmax, h, w, num_samples, &a)) return rc;
    if (!d_lut || !d_order || !d_scratch || !d_queue || !d_emd || !d_status || n_problems < 0 || vmax < 1 ||
        ft_stride < 1)
        return fail(SLB_ERR_INVALID, "slb_emd_solve: bad arguments");
    if (n_problems == 0) return SLB_OK;
    a.lut = d_lut; a.wrap_x = wrap_x; a.wrap_y = wrap_y; a.extra_mass_penalty = extra_mass_penalty;
    const size_t smem = emd_smem_per_warp(vmax) * EMD_WARPS;
    if (smem + kSmemStatic > smem_optin())
        return fail(SLB_ERR_UNSUPPORTED, "slb_emd_solve: %d changed cells in one density map exceed the solver's "
                    "shared memory (%zu bytes per CTA)", vmax, smem);
    if (int rc = allow_smem(emd_solve_kernel, smem)) return rc;
    int dev = 0, sms = 148, per_sm = 1;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    SLB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, emd_solve_kernel, EMD_WARPS * 32, smem));
    long long grid = (n_problems + EMD_WARPS - 1) / EMD_WARPS;
    if (grid > (long long)sms * per_sm) grid = (long long)sms * per_sm;
    const long long by_scratch = (long long)(scratch_doubles / (ft_stride * EMD_WARPS));
    if (by_scratch < 1) return fail(SLB_ERR_INVALID, "slb_emd_solve: scratch smaller than one CTA's flow matrices");
    if (grid > by_scratch) grid = by_scratch;
    emd_solve_kernel<<<(unsigned)grid, EMD_WARPS * 32, smem, (cudaStream_t)stream>>>(
        a, d_order, n_problems, vmax, d_scratch, ft_stride, d_queue, d_emd, d_status);
    SLB_LAUNCHED();
    return SLB_OK;
}

// ---- slb_emd_solve for callers that cannot allocate (no counterpart in the reference) ----------------------
// The solver's scratch — the work-queue word, then the flow matrices — in a file-scope device array that every
// GPU which loads the library holds until exit.  In host code `g_emd_static` is only a stand-in for the symbol:
// the device address comes from cudaGetSymbolAddress, asked once per device with that device current.
constexpr size_t kEmdStaticDoubles = (size_t)SLB_EMD_STATIC_SCRATCH_DOUBLES;
constexpr size_t kEmdStaticHead = 32;  // doubles in front of the flow matrices; the queue word is the first
static __device__ __align__(256) double g_emd_static[kEmdStaticHead + kEmdStaticDoubles];

namespace {
constexpr int kMaxDevices = 64;
std::atomic<double *> g_emd_static_ptr[kMaxDevices];

int emd_static_address(double **out) {
    int dev = 0;
    SLB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(SLB_ERR_UNSUPPORTED, "slb_emd_solve_static: device %d", dev);
    double *p = g_emd_static_ptr[dev].load(std::memory_order_acquire);
    if (!p) {  // (two threads racing here both get the same answer)
        SLB_CUDA(cudaGetSymbolAddress((void **)&p, g_emd_static));
        if (!p) return fail(SLB_ERR_CUDA, "slb_emd_solve_static: cudaGetSymbolAddress returned NULL");
        g_emd_static_ptr[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return SLB_OK;
}
}  // namespace

int slb_emd_solve_static(const uint16_t *d_counts, const uint16_t *d_keys, int n, int kmax, int h, int w,
                         int num_samples, const double *d_lut, int wrap_x, int wrap_y,
                         double extra_mass_penalty, const int32_t *d_order, int n_problems, int vmax,
                         size_t ft_stride, double *d_emd, int32_t *d_status, void *stream) {
    if (ft_stride < 1 || ft_stride * EMD_WARPS > kEmdStaticDoubles)
        return fail(SLB_ERR_UNSUPPORTED,
                    "slb_emd_solve_static: ft_stride = %zu flows per problem exceeds the static scratch "
                    "(%zu doubles, i.e. ft_stride <= %zu); use slb_emd_solve with a scratch of your own",
                    ft_stride, kEmdStaticDoubles, kEmdStaticDoubles / EMD_WARPS);
    double *scratch = nullptr;
    if (int rc = emd_static_address(&scratch)) return rc;
    if (n_problems > 0 && d_counts && d_keys)  // the queue must start at zero; the flow matrices are cleared per problem
        SLB_CUDA(cudaMemsetAsync(scratch, 0, sizeof(uint32_t), (cudaStream_t)stream));
    return slb_emd_solve(d_counts, d_keys, n, kmax, h, w, num_samples, d_lut, wrap_x, wrap_y, extra_mass_penalty,
                         d_order, n_problems, vmax, ft_stride, scratch + kEmdStaticHead, kEmdStaticDoubles,
                         reinterpret_cast<uint32_t *>(scratch), d_emd, d_status, stream);
}

int slb_microbench_int(int kind, int ctas, int iters, uint32_t *d_sink, uint64_t *h_warp_insts, void *stream) {
    if (ctas < 1 || iters < 1 || !d_sink || (kind != 0 && kind != 1))
        return fail(SLB_ERR_INVALID, "slb_microbench_int: bad arguments");
    if (kind == 0) slb::int_peak_kernel<0><<<ctas, 256, 0, (cudaStream_t)stream>>>(d_sink, iters);
    else slb::int_peak_kernel<1><<<ctas, 256, 0, (cudaStream_t)stream>>>(d_sink, iters);
    if (h_warp_insts) *h_warp_insts = (uint64_t)ctas * 8ull * (uint64_t)iters * 8ull;
    SLB_LAUNCHED();
    return SLB_OK;
}

int slb_accumulate_stats(int n, const slb_env_state *d_state, const int32_t *d_reward,
                         const uint8_t *d_done, int64_t *d_stats, void *stream) {
    if (n < 0 || !d_state || !d_stats) return fail(SLB_ERR_INVALID, "slb_accumulate_stats: bad arguments");
    if (n == 0) return SLB_OK;
    int blocks = (n + 255) / 256;
    if (blocks > 148) blocks = 148;
    stats_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(n, d_state, d_reward, d_done,
                                                          (unsigned long long *)d_stats);
    SLB_LAUNCHED();
    return SLB_OK;
}

int slb_get_obs(const slb_pool *pool, int n, const void *d_board, const void *d_goals,
                const slb_env_state *d_state, int vh, int vw, int remove_white_goals,
                uint16_t *d_obs, void *stream) {
    if (int rc = check_pool(pool)) return rc;
    if (n < 0 || vh < 1 || vw < 1 || !d_board || !d_goals || !d_state || !d_obs)
        return fail(SLB_ERR_INVALID, "slb_get_obs: bad arguments");
    if (n == 0) return SLB_OK;
    if (fast_shape(pool->h, pool->w))
        res_obs_kernel<false, true><<<RES_GRID(n), (cudaStream_t)stream>>>(
            *pool, n, (const uint32_t *)d_board, (const uint32_t *)d_goals, d_state, vh, vw,
            remove_white_goals, ObsChannels{}, 0, d_obs);
    else if (big_shape(pool->h, pool->w))
        big_obs_kernel<<<n, 256, 0, (cudaStream_t)stream>>>(*pool, (const uint32_t *)d_board,
                                                          (const uint32_t *)d_goals, d_state, vh, vw,
                                                          remove_white_goals, d_obs);
    else
        get_obs_kernel<<<n, GEN_THREADS, 0, (cudaStream_t)stream>>>(
            *pool, (const uint16_t *)d_board, (const uint16_t *)d_goals, d_state, vh, vw,
            remove_white_goals, d_obs);
    SLB_LAUNCHED();
    return SLB_OK;
}

int slb_side_effect_penalty(const slb_pool *pool, int n, const void *d_board, const void *d_goals,
                            const slb_env_state *d_state, int32_t *d_count, void *stream) {
    if (int rc = check_pool(pool)) return rc;
    if (n < 0 || !d_board || !d_goals || !d_state || !d_count)
        return fail(SLB_ERR_INVALID, "slb_side_effect_penalty: bad arguments");
    if (n == 0) return SLB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (fast_shape(pool->h, pool->w))
        res_penalty_kernel<<<RES_GRID(n), st>>>(*pool, n, (const uint32_t *)d_board, (const uint32_t *)d_goals,
                                               d_state, d_count);
    else if (big_shape(pool->h, pool->w))
        big_penalty_kernel<<<n, 256, 0, st>>>(*pool, (const uint32_t *)d_board, (const uint32_t *)d_goals,
                                             d_state, d_count);
    else
        penalty_generic_kernel<<<n, GEN_THREADS, 0, st>>>(*pool, (const uint16_t *)d_board,
                                                         (const uint16_t *)d_goals, d_state, d_count);
    SLB_LAUNCHED();
    return SLB_OK;
}

static int obs_table(const uint8_t *h_channels, int n_channels, ObsChannels *t, bool *ident) {
    if (!h_channels || n_channels < 1 || n_channels > 16)
        return fail(SLB_ERR_INVALID, "observation channels: 1..16 bit indices expected");
    *t = ObsChannels{};
    *ident = true;
    for (int c = 0; c < n_channels; c++) {
        if (h_channels[c] > 15) return fail(SLB_ERR_INVALID, "observation channel %d is not a bit index", c);
        t->ch[c] = h_channels[c];
        *ident = *ident && h_channels[c] == c;
    }
    return SLB_OK;
}

int slb_get_obs_channels(const slb_pool *pool, int n, const void *d_board, const void *d_goals,
                         const slb_env_state *d_state, int vh, int vw, int remove_white_goals,
                         const uint8_t *h_channels, int n_channels, uint16_t *d_words_scratch,
                         uint16_t *d_out, void *stream) {
    if (int rc = check_pool(pool)) return rc;
    if (n < 0 || vh < 1 || vw < 1 || !d_board || !d_goals || !d_state || !d_out)
        return fail(SLB_ERR_INVALID, "slb_get_obs_channels: bad arguments");
    ObsChannels t;
    bool ident;
    if (int rc = obs_table(h_channels, n_channels, &t, &ident)) return rc;
    if ((uintptr_t)d_out % 16) return fail(SLB_ERR_INVALID, "slb_get_obs_channels: d_out must be 16-byte aligned");
    if (n == 0) return SLB_OK;
    if (fast_shape(pool->h, pool->w) && (long long)vh * vw <= RES_VIEW_U16) {
        if (ident)
            res_obs_kernel<true, true><<<RES_GRID(n), (cudaStream_t)stream>>>(
                *pool, n, (const uint32_t *)d_board, (const uint32_t *)d_goals, d_state, vh, vw,
                remove_white_goals, t, n_channels, d_out);
        else
            res_obs_kernel<true, false><<<RES_GRID(n), (cudaStream_t)stream>>>(
                *pool, n, (const uint32_t *)d_board, (const uint32_t *)d_goals, d_state, vh, vw,
                remove_white_goals, t, n_channels, d_out);
        SLB_LAUNCHED();
        return SLB_OK;
    }
    if (!d_words_scratch)
        return fail(SLB_ERR_UNSUPPORTED, "slb_get_obs_channels: this board / view shape needs d_words_scratch");
    if (int rc = slb_get_obs(pool, n, d_board, d_goals, d_state, vh, vw, remove_white_goals, d_words_scratch, stream))
        return rc;
    return slb_obs_channels(d_words_scratch, (size_t)n * vh * vw, h_channels, n_channels, d_out, stream);
}

static int env_step_obs_impl(const slb_pool *pool, int n, void *d_board, void *d_goals, slb_env_state *d_state,
                             const uint8_t *d_actions, int time_limit, uint64_t seed, uint64_t env_id0,
                             int32_t *d_reward, uint8_t *d_done, int auto_reset_stride, int vh, int vw,
                             int remove_white_goals, const uint8_t *h_channels, int n_channels, uint16_t *d_obs,
                             void *stream, unsigned long long *gc) {
    if (int rc = check_pool(pool)) return rc;
    if (n < 0 || vh < 1 || vw < 1 || !d_board || !d_goals || !d_state || !d_actions || !d_obs)
        return fail(SLB_ERR_INVALID, "slb_env_step_obs: bad arguments");
    ObsArgs oa = {};
    bool ident = true;
    if (n_channels != 0) {
        if (int rc = obs_table(h_channels, n_channels, &oa.tab, &ident)) return rc;
        if ((uintptr_t)d_obs % 16) return fail(SLB_ERR_INVALID, "slb_env_step_obs: d_obs must be 16-byte aligned");
    }
    if (!fast_shape(pool->h, pool->w) || (n_channels != 0 && (long long)vh * vw > RES_VIEW_U16))
        return fail(SLB_ERR_UNSUPPORTED, "slb_env_step_obs: this board / view shape takes slb_env_step + slb_get_obs");
    if ((uint64_t)n * 16u * (uint64_t)pool->h >= (1ull << 32))
        return fail(SLB_ERR_UNSUPPORTED, "too many envs per launch (resident word offsets are 32-bit)");
    if (n == 0) return SLB_OK;
    oa.vh = vh; oa.vw = vw; oa.remove_white = remove_white_goals; oa.c_n = n_channels; oa.obs = d_obs;
    SLB_CUDA(launch_res_env_step_obs(*pool, n, (uint32_t *)d_board, (uint32_t *)d_goals, d_state, d_actions,
                                     time_limit, seed, env_id0, d_reward, d_done, auto_reset_stride, oa, ident,
                                     (cudaStream_t)stream, gc));
    SLB_LAUNCHED();
    return SLB_OK;
}

int slb_env_step_obs(const slb_pool *pool, int n, void *d_board, void *d_goals, slb_env_state *d_state,
                     const uint8_t *d_actions, int time_limit, uint64_t seed, uint64_t env_id0,
                     int32_t *d_reward, uint8_t *d_done, int auto_reset_stride, int vh, int vw,
                     int remove_white_goals, const uint8_t *h_channels, int n_channels, uint16_t *d_obs,
                     void *stream) {
    return env_step_obs_impl(pool, n, d_board, d_goals, d_state, d_actions, time_limit, seed, env_id0, d_reward,
                             d_done, auto_reset_stride, vh, vw, remove_white_goals, h_channels, n_channels, d_obs,
                             stream, nullptr);
}

// ---- device-buffer step with every per-batch argument bound once (the GPU-policy loop) -----------
struct slb_dev_stepper {
    const slb_pool *pool;
    int n, time_limit, stride;
    void *board, *goals;
    slb_env_state *state;
    uint64_t seed, env_id0;
    int32_t *reward;
    int vh, vw, remove_white, n_channels;  // vh == 0: no observation
    uint8_t channels[16];
    unsigned long long *gc;  // SafeLifeEnv.global_counter on the device (may be NULL)
    uint8_t *already;        // [n] scratch for the counter on shapes whose step kernel does not count
};

int slb_dev_stepper_create(const slb_pool *pool, int n, void *d_board, void *d_goals, slb_env_state *d_state,
                           int time_limit, uint64_t seed, uint64_t env_id0, int auto_reset_stride,
                           int32_t *d_reward, int vh, int vw, int remove_white_goals,
                           const uint8_t *h_channels, int n_channels, uint64_t *d_global_counter,
                           slb_dev_stepper **out) {
    if (int rc = check_pool(pool)) return rc;
    if (!out || n < 0 || !d_board || !d_goals || !d_state || vh < 0 || vw < 0 || n_channels < 0 || n_channels > 16 ||
        (n_channels > 0 && !h_channels))
        return fail(SLB_ERR_INVALID, "slb_dev_stepper_create: bad arguments");
    slb_dev_stepper *s = new slb_dev_stepper{pool, n, time_limit, auto_reset_stride, d_board, d_goals, d_state,
                                             seed, env_id0, d_reward, vh, vw, remove_white_goals, n_channels, {},
                                             (unsigned long long *)d_global_counter, nullptr};
    for (int c = 0; c < n_channels; c++) s->channels[c] = h_channels[c];
    if (d_global_counter && !fast_shape(pool->h, pool->w) && n > 0) {
        if (cudaMalloc(&s->already, (size_t)n) != cudaSuccess) {
            delete s;
            return fail(SLB_ERR_CUDA, "slb_dev_stepper_create: cudaMalloc failed");
        }
    }
    *out = s;
    return SLB_OK;
}

int slb_dev_stepper_step(slb_dev_stepper *s, const uint8_t *d_actions, int32_t *d_reward, uint8_t *d_done,
                         uint16_t *d_obs, void *stream) {
    if (!s) return fail(SLB_ERR_INVALID, "stepper is NULL");
    int32_t *reward = d_reward ? d_reward : s->reward;
    if (s->vh > 0 && d_obs)
        return env_step_obs_impl(s->pool, s->n, s->board, s->goals, s->state, d_actions, s->time_limit, s->seed,
                                 s->env_id0, reward, d_done, s->stride, s->vh, s->vw, s->remove_white,
                                 s->n_channels ? s->channels : nullptr, s->n_channels, d_obs, stream, s->gc);
    if (s->gc && s->already) {
        // shapes whose step kernel does not count: who was already done, then the step, then the counts
        if (!d_done) return fail(SLB_ERR_INVALID, "slb_dev_stepper_step: the global counter needs d_done");
        gc_pre_kernel<<<(s->n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(s->n, s->state, s->already);
        SLB_LAUNCHED();
        if (int rc = env_step_impl(s->pool, s->n, s->board, s->goals, s->state, d_actions, s->time_limit, s->seed,
                                   s->env_id0, reward, d_done, s->stride, stream, nullptr)) return rc;
        gc_post_kernel<<<(s->n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(s->n, d_done, s->already, s->gc);
        SLB_LAUNCHED();
        return SLB_OK;
    }
    return env_step_impl(s->pool, s->n, s->board, s->goals, s->state, d_actions, s->time_limit, s->seed,
                         s->env_id0, reward, d_done, s->stride, stream, s->gc);
}

void slb_dev_stepper_destroy(slb_dev_stepper *s) {
    if (!s) return;
    if (s->already) cudaFree(s->already);
    delete s;
}

int slb_obs_channels(const uint16_t *d_words, size_t m, const uint8_t *h_channels, int n_channels,
                     uint16_t *d_out, void *stream) {
    if (!d_words || !d_out) return fail(SLB_ERR_INVALID, "slb_obs_channels: bad arguments");
    if ((uintptr_t)d_out % 16) return fail(SLB_ERR_INVALID, "slb_obs_channels: d_out must be 16-byte aligned");
    ObsChannels t;
    bool ident;
    if (int rc = obs_table(h_channels, n_channels, &t, &ident)) return rc;
    if (m == 0) return SLB_OK;
    size_t chunks = (m * (size_t)n_channels) >> 3;
    size_t blocks = (chunks + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > 148 * 32) blocks = 148 * 32;
    if (ident)
        obs_channels_kernel<true><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_words, m, t, n_channels, d_out);
    else
        obs_channels_kernel<false><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_words, m, t, n_channels, d_out);
    SLB_LAUNCHED();
    return SLB_OK;
}

}  // extern "C"
