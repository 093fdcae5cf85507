/*
 * stream_random.c — ORACLE-SIDE shim (test infrastructure only).
 *
 * Replacement for the reference's speedups_src/random.c (which pulls doubles from the global
 * numpy RNG through the CPython API, random.c:14-26,47-52) so that the reference's own
 * advance_board.c — compiled unchanged, in place — can be linked WITHOUT Python and fed the
 * build's counter-based per-env stream (second alternative).  It implements the
 * three functions the reference's random.h declares; advance_board.c only references
 * random_float() (advance_board.c:4,110).
 *
 * The stream is thread-local, so one OpenMP/ctypes thread per env slice is safe.
 */
#include <stdint.h>

static _Thread_local uint64_t t_seed, t_env, t_counter;

void ref_stream_set(uint64_t seed, uint64_t env, uint64_t counter) {
    t_seed = seed;
    t_env = env;
    t_counter = counter;
}

uint64_t ref_stream_get_counter(void) { return t_counter; }

/* Philox4x32-10, third independent statement (oracle C, numpy feeder and CUDA are the others) */
static void philox(uint32_t c[4], uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; r++) {
        uint64_t a = 0xD2511F53ull * c[0], b = 0xCD9E8D57ull * c[2];
        uint32_t t0 = (uint32_t)(b >> 32) ^ c[1] ^ k0, t2 = (uint32_t)(a >> 32) ^ c[3] ^ k1;
        c[1] = (uint32_t)b;
        c[3] = (uint32_t)a;
        c[0] = t0;
        c[2] = t2;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
}

double random_float(void) {
    uint32_t c[4] = {(uint32_t)t_counter, (uint32_t)(t_counter >> 32), (uint32_t)t_env,
                     (uint32_t)(t_env >> 32)};
    philox(c, (uint32_t)t_seed, (uint32_t)(t_seed >> 32));
    t_counter++;
    uint64_t u = ((uint64_t)(c[0] >> 5) << 26) | (uint64_t)(c[1] >> 6);
    return (double)u / 9007199254740992.0;
}

int32_t random_int(int32_t high) { return (int32_t)(random_float() * high); }

int random_seed(uint32_t seed) {
    t_seed = seed;
    t_counter = 0;
    return 0;
}
