// ctx.cu -- context, error plumbing, scratch pools and the host-side R-compatible RNG (f1).
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

int ate_fail(ate_ctx *ctx, int code, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf;
    return code;
}

static thread_local std::string g_create_err;

extern "C" int ate_version(void) { return 100; }

extern "C" int ate_ctx_create(int device, ate_ctx **out) {
    if (!out) return -1;
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0) {
        // no CPU fallback by design: the caller must see this
        g_create_err = std::string("no CUDA device: ") + cudaGetErrorString(e);
        fprintf(stderr, "libate_b200: %s\n", g_create_err.c_str());
        return -2;
    }
    if (device < 0 || device >= ndev) return -3;
    if (cudaSetDevice(device) != cudaSuccess) return -4;
    ate_ctx *c = new ate_ctx();
    c->device = device;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) c->num_sms = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete c;
        return -5;
    }
    c->rng.seeded = false;
    *out = c;
    return 0;
}

extern "C" void ate_ctx_destroy(ate_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    ate_ctx_nccl_destroy(ctx);
    for (auto &pr : ctx->prof_events) {
        cudaEventDestroy(pr.first);
        cudaEventDestroy(pr.second);
    }
    for (auto &kv : ctx->cache)
        if (kv.second.first) cudaFree(kv.second.first);
    for (auto &kv : ctx->pinned)
        if (kv.second.first) cudaFreeHost(kv.second.first);
    if (ctx->scratch) cudaFree(ctx->scratch);
    if (ctx->gram_scratch) cudaFree(ctx->gram_scratch);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" const char *ate_last_error(const ate_ctx *ctx) {
    return ctx ? ctx->err.c_str() : g_create_err.c_str();
}

extern "C" void *ate_ctx_stream(const ate_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }

extern "C" int ate_ctx_set_allreduce(ate_ctx *ctx, ate_allreduce_fn fn, void *user) {
    if (!ctx) return -1;
    ctx->allreduce = fn;
    ctx->allreduce_user = user;
    return 0;
}

extern "C" long long ate_launch_count(const ate_ctx *ctx) { return ctx ? ctx->launches : 0; }

extern "C" int ate_profile_enable(ate_ctx *ctx, int on) {
    if (!ctx) return -1;
    cudaStreamSynchronize(ctx->stream);
    for (auto &pr : ctx->prof_events) {
        cudaEventDestroy(pr.first);
        cudaEventDestroy(pr.second);
    }
    ctx->prof_events.clear();
    ctx->prof_shape.clear();
    ctx->prof_flops = ctx->prof_bytes = 0.0;
    ctx->prof_launches = 0;
    ctx->profile = on < 0 ? 0 : (on & 3);
    return 0;
}

extern "C" int ate_profile_read(ate_ctx *ctx, double *gram_ms, long long *gram_launches,
                                double *gram_flops, double *gram_bytes) {
    if (!ctx) return -1;
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    double ms = 0.0;
    // ATE_PROFILE_CSV=<file>: one line per timed launch (columns in the Gram, rows, ms)
    const char *csv = getenv("ATE_PROFILE_CSV");
    FILE *fp = csv ? fopen(csv, "w") : nullptr;
    if (fp) fprintf(fp, "launch,gram_cols,rows,ms\n");
    size_t k = 0;
    for (auto &pr : ctx->prof_events) {
        float t = 0.f;
        ATE_CUDA(ctx, cudaEventElapsedTime(&t, pr.first, pr.second));
        ms += t;
        if (fp && k < ctx->prof_shape.size())
            fprintf(fp, "%zu,%d,%lld,%.6f\n", k, ctx->prof_shape[k].first, ctx->prof_shape[k].second, (double)t);
        ++k;
    }
    if (fp) fclose(fp);
    if (gram_ms) *gram_ms = ms;
    if (gram_launches) *gram_launches = ctx->prof_launches;
    if (gram_flops) *gram_flops = ctx->prof_flops;
    if (gram_bytes) *gram_bytes = ctx->prof_bytes;
    return 0;
}

static int grow(ate_ctx *ctx, void **buf, size_t *have, size_t want, void **out) {
    if (want > *have) {
        ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (*buf) ATE_CUDA(ctx, cudaFree(*buf));
        *buf = nullptr;
        *have = 0;
        size_t cap = want + (want >> 2) + 4096;
        ATE_CUDA(ctx, cudaMalloc(buf, cap));
        if (getenv("ATE_POISON")) {
            ATE_CUDA(ctx, cudaMemset(*buf, 0xFF, cap));
            ATE_CUDA(ctx, cudaDeviceSynchronize());
        }
        *have = cap;
    }
    *out = *buf;
    return 0;
}

int ate_cached(ate_ctx *ctx, const char *tag, size_t bytes, void **out) {
    auto &slot = ctx->cache[tag];
    return grow(ctx, &slot.first, &slot.second, bytes, out);
}

int ate_cached_pinned(ate_ctx *ctx, const char *tag, size_t bytes, void **out) {
    auto &slot = ctx->pinned[tag];
    if (bytes > slot.second) {
        ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (slot.first) ATE_CUDA(ctx, cudaFreeHost(slot.first));
        slot.first = nullptr;
        slot.second = 0;
        const size_t cap = bytes + (bytes >> 2) + 4096;
        ATE_CUDA(ctx, cudaHostAlloc(&slot.first, cap, cudaHostAllocDefault));
        slot.second = cap;
    }
    *out = slot.first;
    return 0;
}

int ate_scratch(ate_ctx *ctx, size_t bytes, void **out) {
    return grow(ctx, &ctx->scratch, &ctx->scratch_bytes, bytes, out);
}
int ate_gram_scratch(ate_ctx *ctx, size_t bytes, void **out) {
    return grow(ctx, &ctx->gram_scratch, &ctx->gram_scratch_bytes, bytes, out);
}

// ---- sd() as R computes it: two passes, n-1 --------------------------------------------------
double ate_host_sd(const double *x, long long n) {
    if (n < 2) return NAN;
    long double s = 0.0L;
    for (long long i = 0; i < n; ++i) s += x[i];
    long double mu = s / (long double)n, t = 0.0L;
    for (long long i = 0; i < n; ++i) t += (x[i] - mu);
    mu += t / (long double)n;
    double m = (double)mu;
    long double ss = 0.0L;
    for (long long i = 0; i < n; ++i) {
        long double d = (long double)x[i] - m;
        ss += d * d;
    }
    return sqrt((double)(ss / (long double)(n - 1)));
}

// ---- R's Mersenne-Twister stream: set.seed(), unif_rand(), sample() -------------------------
// Initial scrambling is the LCG 69069*s+1; 50 warm-up steps, one more for the slot that R
// then overwrites with the position counter, then 624 state words.
void ate_host_rng_seed(ate_rng_state *r, uint32_t s) {
    for (int i = 0; i < 51; ++i) s = s * 69069u + 1u;
    for (int i = 0; i < 624; ++i) {
        s = s * 69069u + 1u;
        r->mt[i] = s;
    }
    r->mti = 624;  // forces a twist before the first output
    r->seeded = true;
}

static inline uint32_t mt_draw(ate_rng_state *r) {
    if (r->mti >= 624) {
        uint32_t *m = r->mt;
        for (int i = 0; i < 624; ++i) {
            uint32_t y = (m[i] & 0x80000000u) | (m[(i + 1) % 624] & 0x7fffffffu);
            uint32_t z = m[(i + 397) % 624] ^ (y >> 1);
            if (y & 1u) z ^= 0x9908b0dfu;
            m[i] = z;
        }
        r->mti = 0;
    }
    uint32_t y = r->mt[r->mti++];
    y ^= y >> 11;
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= y >> 18;
    return y;
}

static inline double r_unif(ate_rng_state *r) {
    double u = (double)mt_draw(r) * 2.3283064365386963e-10;
    const double half_ulp = 0.5 * 2.328306437080797e-10;
    if (u <= 0.0) return half_ulp;
    if (1.0 - u <= 0.0) return 1.0 - half_ulp;
    return u;
}

static inline double r_index(ate_rng_state *r, double dn, int kind) {
    if (kind == ATE_RNG_ROUNDING) return floor(dn * r_unif(r));
    if (dn <= 0.0) return 0.0;
    const int bits = (int)ceil(log2(dn));
    for (;;) {
        int64_t acc = 0;
        for (int got = 0; got <= bits; got += 16) acc = acc * 65536 + (int)floor(r_unif(r) * 65536.0);
        if (bits < 64) acc &= (((int64_t)1) << bits) - 1;
        if ((double)acc < dn) return (double)acc;
    }
}

void ate_host_sample_replace(ate_rng_state *r, int n, long long k, int kind, int32_t *out) {
    for (long long i = 0; i < k; ++i) out[i] = (int32_t)r_index(r, (double)n, kind) + 1;
}

void ate_host_foldid(ate_rng_state *r, int n, int nfolds, int kind, int32_t *out) {
    // sample(v) == v[sample.int(n)], v = rep(seq(nfolds), length = n)
    std::vector<int32_t> pool((size_t)n);
    for (int i = 0; i < n; ++i) pool[i] = i;
    int left = n;
    for (int i = 0; i < n; ++i) {
        int j = (int)r_index(r, (double)left, kind);
        out[i] = pool[j] % nfolds + 1;
        pool[j] = pool[--left];
    }
}

extern "C" int ate_r_set_seed(ate_ctx *ctx, uint32_t seed) {
    if (!ctx) return -1;
    ate_host_rng_seed(&ctx->rng, seed);
    return 0;
}

extern "C" int ate_r_sample(ate_ctx *ctx, int n, long long k, int rng_kind, int32_t *out) {
    if (!ctx || !out || n <= 0 || k < 0) return ate_fail(ctx, -1, "ate_r_sample: bad arguments");
    if (!ctx->rng.seeded) return ate_fail(ctx, -1, "ate_r_sample: call ate_r_set_seed first");
    ate_host_sample_replace(&ctx->rng, n, k, rng_kind, out);
    return 0;
}

extern "C" int ate_r_foldid(ate_ctx *ctx, int n, int nfolds, int rng_kind, int32_t *out) {
    if (!ctx || !out || n <= 0 || nfolds <= 0)
        return ate_fail(ctx, -1, "ate_r_foldid: bad arguments");
    if (!ctx->rng.seeded) return ate_fail(ctx, -1, "ate_r_foldid: call ate_r_set_seed first");
    ate_host_foldid(&ctx->rng, n, nfolds, rng_kind, out);
    return 0;
}
