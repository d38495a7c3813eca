// Host-side replica of torch's CPU generator streams (mt19937 + ATen's distribution transforms), so that the
// random numbers the reference consumes inside every Adam iteration of Actor.query can be produced ahead of time,
// in bulk, in exactly the reference's order and bit for bit:
//   amortized_network.py:136      torch.rand(N)            float32, one mt19937 word per element
//   amortized_network.py:156      N x torch.randperm(d)    Fisher-Yates, one word per swap
//   acqfs.py:624                  torch.randn_like(diff)   float64, ATen normal_fill (N >= 16) / serial Box-Muller
// ("same inputs and seeds => identical next query").  The state blob is the 5056-byte tensor of
// torch.get_rng_state() / set_rng_state() (ATen CPUGeneratorImplState); the functions advance it in place.
// Pure host code: no CUDA calls.  glibc's log / sin / cos / sqrt are the functions ATen calls for float64.
#include <math.h>
#include <stdint.h>
#include <string.h>
#include <thread>
#include <vector>
#include "common.cuh"

namespace nmb {
namespace {

constexpr int kMtN = 624, kMtM = 397;
constexpr size_t kBlobBytes = 5056;
// offsets inside CPUGeneratorImplState (legacy POD first)
constexpr size_t kOffLeft = 8, kOffSeeded = 12, kOffNext = 16, kOffState = 24;
constexpr size_t kOffNormalY = kOffState + 8 * kMtN + 8, kOffNormalValid = kOffNormalY + 16;

struct Mt19937 {
    uint32_t st[kMtN];
    int left, next;
    bool has_cached;      // next_double_normal_sample_
    double cached;

    void load(const uint8_t* blob) {
        int32_t l, s;
        uint64_t nx;
        memcpy(&l, blob + kOffLeft, 4);
        memcpy(&s, blob + kOffSeeded, 4);
        memcpy(&nx, blob + kOffNext, 8);
        left = l;
        next = (int)nx;
        for (int i = 0; i < kMtN; ++i) {
            uint64_t v;
            memcpy(&v, blob + kOffState + 8 * i, 8);
            st[i] = (uint32_t)v;
        }
        int32_t valid;
        memcpy(&valid, blob + kOffNormalValid, 4);
        has_cached = valid != 0;
        memcpy(&cached, blob + kOffNormalY, 8);
    }
    void store(uint8_t* blob) const {
        int32_t l = left;
        uint64_t nx = (uint64_t)next;
        memcpy(blob + kOffLeft, &l, 4);
        memcpy(blob + kOffNext, &nx, 8);
        for (int i = 0; i < kMtN; ++i) {
            uint64_t v = st[i];
            memcpy(blob + kOffState + 8 * i, &v, 8);
        }
        int32_t valid = has_cached ? 1 : 0;
        double y = has_cached ? cached : 0.0;
        memcpy(blob + kOffNormalValid, &valid, 4);
        memcpy(blob + kOffNormalY, &y, 8);
    }
    static uint32_t mix(uint32_t u, uint32_t v) { return (u & 0x80000000u) | (v & 0x7fffffffu); }
    static uint32_t twist(uint32_t u, uint32_t v) { return (mix(u, v) >> 1) ^ ((v & 1u) ? 0x9908b0dfu : 0u); }
    void reload() {
        left = kMtN;
        next = 0;
        // index loops with a dependency distance of 227 words: the compiler vectorises them
        for (int i = 0; i < kMtN - kMtM; ++i) st[i] = st[i + kMtM] ^ twist(st[i], st[i + 1]);
        for (int i = kMtN - kMtM; i < kMtN - 1; ++i) st[i] = st[i + kMtM - kMtN] ^ twist(st[i], st[i + 1]);
        st[kMtN - 1] = st[kMtM - 1] ^ twist(st[kMtN - 1], st[0]);
    }
    uint32_t u32() {
        if (--left == 0) reload();
        uint32_t y = st[next++];
        y ^= (y >> 11);
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= (y >> 18);
        return y;
    }
    // the next n tempered words, exactly as n calls of u32() (bulk: the tempering loop vectorises)
    void words(uint32_t* out, size_t n) {
        while (n) {
            if (left == 1) {
                reload();
                left = kMtN + 1;      // nothing consumed yet: left - 1 words available at next = 0
            }
            const size_t avail = (size_t)left - 1, take = n < avail ? n : avail;
            const uint32_t* src = st + next;
            for (size_t i = 0; i < take; ++i) {
                uint32_t y = src[i];
                y ^= (y >> 11);
                y ^= (y << 7) & 0x9d2c5680u;
                y ^= (y << 15) & 0xefc60000u;
                y ^= (y >> 18);
                out[i] = y;
            }
            next += (int)take;
            left -= (int)take;
            out += take;
            n -= take;
        }
    }
    uint64_t u64() {
        const uint32_t hi = u32(), lo = u32();
        return ((uint64_t)hi << 32) | lo;
    }
    // at::uniform_real_distribution<float>(0, 1): 24 bits of one word
    float uniform_f32() { return (float)(u32() & ((1u << 24) - 1)) * (1.0f / 16777216.0f); }
    // at::uniform_real_distribution<double>(0, 1): 53 bits of two words
    double uniform_f64() { return (double)(u64() & ((1ULL << 53) - 1)) * (1.0 / 9007199254740992.0); }
};

// ATen normal_fill_16<double>(data, 0, 1)
inline void normal_fill_16(double* data) {
    for (int j = 0; j < 8; ++j) {
        const double u1 = 1 - data[j];
        const double u2 = data[j + 8];
        const double radius = sqrt(-2 * log(u1));
        const double theta = 2.0f * 3.14159265358979323846 * u2;
        data[j] = radius * cos(theta) * 1.0 + 0.0;
        data[j + 8] = radius * sin(theta) * 1.0 + 0.0;
    }
}

// the word-consuming half of torch.randn(n, dtype=float64) on the CPU generator; n >= 16 leaves uniforms behind that
// finish_normal() turns into normals (order-independent, so many calls can be finished in parallel)
inline void randn_f64_draw(Mt19937& g, int64_t n, double* out, double* tail_redraw /* 16 or null */) {
    if (n >= 16) {
        for (int64_t i = 0; i < n; ++i) out[i] = g.uniform_f64();
        if (n % 16 != 0)
            for (int i = 0; i < 16; ++i) tail_redraw[i] = g.uniform_f64();
        return;
    }
    // serial at::normal_distribution<double>: Box-Muller with the second value cached in the generator
    for (int64_t i = 0; i < n; ++i) {
        if (g.has_cached) {
            out[i] = g.cached;
            g.has_cached = false;
            continue;
        }
        const double u1 = g.uniform_f64();
        const double u2 = g.uniform_f64();
        const double r = sqrt(-2.0 * log1p(-u2));
        const double theta = 2.0 * 3.14159265358979323846 * u1;
        g.cached = r * sin(theta);
        g.has_cached = true;
        out[i] = r * cos(theta);
    }
}

inline void randn_f64_finish(int64_t n, double* out, const double* tail_redraw) {
    if (n < 16) return;
    for (int64_t i = 0; i + 15 < n; i += 16) normal_fill_16(out + i);
    if (n % 16 != 0) {
        double* last = out + n - 16;
        memcpy(last, tail_redraw, 16 * sizeof(double));
        normal_fill_16(last);
    }
}

// one noisy policy call (amortized_network.py:136,156): rand(N) then N x randperm(d)[:2]
inline void rotation_draw(Mt19937& g, int64_t N, int d, float* u, int32_t* dims) {
    for (int64_t i = 0; i < N; ++i) u[i] = g.uniform_f32();
    int64_t perm[64];
    for (int64_t i = 0; i < N; ++i) {
        for (int k = 0; k < d; ++k) perm[k] = k;
        for (int k = 0; k < d - 1; ++k) {
            const int64_t z = (int64_t)(g.u32() % (uint32_t)(d - k));
            const int64_t sav = perm[k];
            perm[k] = perm[z + k];
            perm[z + k] = sav;
        }
        dims[2 * i] = (int32_t)perm[0];
        dims[2 * i + 1] = (int32_t)(d > 1 ? perm[1] : perm[0]);
    }
}

// ---- the same transforms over a pre-generated word buffer (positions are known in advance when N >= 16)
struct Words {
    const uint32_t* p;
    uint32_t u32() { return *p++; }
    float uniform_f32() { return (float)(u32() & ((1u << 24) - 1)) * (1.0f / 16777216.0f); }
    double uniform_f64() {
        const uint32_t hi = u32(), lo = u32();
        return (double)((((uint64_t)hi << 32) | lo) & ((1ULL << 53) - 1)) * (1.0 / 9007199254740992.0);
    }
};

inline void rotation_from_words(Words& w, int64_t N, int d, float* u, int32_t* dims) {
    for (int64_t i = 0; i < N; ++i) u[i] = w.uniform_f32();
    int64_t perm[64];
    for (int64_t i = 0; i < N; ++i) {
        for (int k = 0; k < d; ++k) perm[k] = k;
        for (int k = 0; k < d - 1; ++k) {
            const int64_t z = (int64_t)(w.u32() % (uint32_t)(d - k));
            const int64_t sav = perm[k];
            perm[k] = perm[z + k];
            perm[z + k] = sav;
        }
        dims[2 * i] = (int32_t)perm[0];
        dims[2 * i + 1] = (int32_t)(d > 1 ? perm[1] : perm[0]);
    }
}

inline void randn_from_words(Words& w, int64_t N, double* out) {      // N >= 16
    for (int64_t i = 0; i < N; ++i) out[i] = w.uniform_f64();
    double tail[16];
    if (N % 16 != 0)
        for (int i = 0; i < 16; ++i) tail[i] = w.uniform_f64();
    randn_f64_finish(N, out, tail);
}

}  // namespace
}  // namespace nmb

using namespace nmb;

extern "C" int64_t nmb_host_rng_state_bytes(void) { return (int64_t)kBlobBytes; }

extern "C" int nmb_host_rng_rand_f32(uint8_t* state, int64_t n, float* out) {
    NMB_REQUIRE(state && (out || n == 0) && n >= 0, NMB_EINVAL, "nmb_host_rng_rand_f32: null argument");
    Mt19937 g;
    g.load(state);
    for (int64_t i = 0; i < n; ++i) out[i] = g.uniform_f32();
    g.store(state);
    return 0;
}

extern "C" int nmb_host_rng_randperm(uint8_t* state, int n_perms, int d, int64_t* out) {
    NMB_REQUIRE(state && out && n_perms >= 0 && d >= 1, NMB_EINVAL, "nmb_host_rng_randperm: bad argument");
    Mt19937 g;
    g.load(state);
    for (int p = 0; p < n_perms; ++p) {
        int64_t* r = out + (size_t)p * d;
        for (int k = 0; k < d; ++k) r[k] = k;
        for (int k = 0; k < d - 1; ++k) {
            const int64_t z = (int64_t)(g.u32() % (uint32_t)(d - k));
            const int64_t sav = r[k];
            r[k] = r[z + k];
            r[z + k] = sav;
        }
    }
    g.store(state);
    return 0;
}

extern "C" int nmb_host_rng_randn_f64(uint8_t* state, int n_calls, const int64_t* sizes, double* out, int n_threads) {
    NMB_REQUIRE(state && sizes && out && n_calls >= 0, NMB_EINVAL, "nmb_host_rng_randn_f64: null argument");
    Mt19937 g;
    g.load(state);
    size_t total = 0;
    bool all_bulk = true;
    for (int c = 0; c < n_calls; ++c) {
        NMB_REQUIRE(sizes[c] >= 0, NMB_EINVAL, "nmb_host_rng_randn_f64: negative size");
        total += (size_t)sizes[c];
        if (sizes[c] > 0 && sizes[c] < 16) all_bulk = false;
    }
    if (all_bulk && total >= 32768 && n_threads > 1) {
        // every call consumes a known number of words: serial word stream, parallel transform in 16-value blocks
        std::vector<size_t> woff(n_calls + 1, 0), ooff(n_calls + 1, 0);
        for (int c = 0; c < n_calls; ++c) {
            woff[c + 1] = woff[c] + 2 * (size_t)sizes[c] + ((sizes[c] % 16) ? 32 : 0);
            ooff[c + 1] = ooff[c] + (size_t)sizes[c];
        }
        std::vector<uint32_t> buf(woff[n_calls]);
        g.words(buf.data(), buf.size());
        g.store(state);
        auto u64at = [&](size_t w) {
            return (double)((((uint64_t)buf[w] << 32) | buf[w + 1]) & ((1ULL << 53) - 1)) * (1.0 / 9007199254740992.0);
        };
        auto work = [&](int t, int nt) {
            for (int c = 0; c < n_calls; ++c) {
                const int64_t n = sizes[c];
                if (n == 0) continue;
                double* dst = out + ooff[c];
                const int64_t nblk = n / 16;
                for (int64_t b = nblk * t / nt; b < nblk * (t + 1) / nt; ++b) {
                    for (int i = 0; i < 16; ++i) dst[16 * b + i] = u64at(woff[c] + 2 * (size_t)(16 * b + i));
                    normal_fill_16(dst + 16 * b);
                }
            }
        };
        std::vector<std::thread> pool;
        for (int t = 0; t < n_threads; ++t) pool.emplace_back(work, t, n_threads);
        for (auto& th : pool) th.join();
        for (int c = 0; c < n_calls; ++c) {          // ragged tails: the last 16 values are recomputed from fresh uniforms
            const int64_t n = sizes[c];
            if (n == 0 || n % 16 == 0) continue;
            double* last = out + ooff[c] + n - 16;
            for (int i = 0; i < 16; ++i) last[i] = u64at(woff[c] + 2 * (size_t)n + 2 * (size_t)i);
            normal_fill_16(last);
        }
        return 0;
    }
    double tail[16];
    size_t o = 0;
    for (int c = 0; c < n_calls; ++c) {
        randn_f64_draw(g, sizes[c], out + o, tail);
        randn_f64_finish(sizes[c], out + o, tail);
        o += (size_t)sizes[c];
    }
    g.store(state);
    return 0;
}

extern "C" int nmb_host_rng_ts_am_stream(uint8_t* state, int n_iters, int L, int64_t N, int d, float* u_main,
                                         int32_t* dims_main, double* nz_main, float* u_orig, int32_t* dims_orig,
                                         double* nz_orig, uint8_t* states_out, int n_threads) {
    NMB_REQUIRE(state && u_main && dims_main && nz_main && u_orig && dims_orig && nz_orig, NMB_EINVAL,
                "nmb_host_rng_ts_am_stream: null argument");
    NMB_REQUIRE(n_iters >= 0 && L >= 0 && N >= 1 && d >= 1 && d <= 64, NMB_EINVAL,
                "nmb_host_rng_ts_am_stream: n_iters=%d L=%d N=%lld d=%d", n_iters, L, (long long)N, d);
    Mt19937 g;
    g.load(state);
    const int calls = L + 1;
    const bool bulk = N >= 16, ragged = bulk && (N % 16 != 0);
    if (bulk && !states_out) {
        // fast path: every call consumes a known number of words, so the serial part is only the mt19937 word
        // stream; the transforms run per iteration on `n_threads` host threads
        const size_t w_rot = (size_t)N * (size_t)d, w_nz = (size_t)2 * N + (ragged ? 32 : 0);
        const size_t w_it = (size_t)(1 + calls) * w_rot + (size_t)2 * calls * w_nz;
        std::vector<uint32_t> buf((size_t)n_iters * w_it);
        g.words(buf.data(), buf.size());
        g.store(state);
        auto work = [&](int i0, int i1) {
            for (int it = i0; it < i1; ++it) {
                Words w{buf.data() + (size_t)it * w_it};
                rotation_from_words(w, N, d, u_main + (size_t)it * N, dims_main + (size_t)it * N * 2);
                for (int c = 0; c < calls; ++c) randn_from_words(w, N, nz_main + ((size_t)it * calls + c) * N);
                for (int c = 0; c < calls; ++c)
                    rotation_from_words(w, N, d, u_orig + ((size_t)it * calls + c) * N,
                                        dims_orig + ((size_t)it * calls + c) * N * 2);
                for (int c = 0; c < calls; ++c) randn_from_words(w, N, nz_orig + ((size_t)it * calls + c) * N);
            }
        };
        int nt = n_threads < 1 ? 1 : n_threads;
        if (nt > n_iters) nt = n_iters;
        if (nt <= 1 || (size_t)n_iters * w_it < 200000) {
            work(0, n_iters);
        } else {
            std::vector<std::thread> pool;
            for (int t = 0; t < nt; ++t)
                pool.emplace_back(work, (int)((long long)n_iters * t / nt), (int)((long long)n_iters * (t + 1) / nt));
            for (auto& th : pool) th.join();
        }
        return 0;
    }
    // tail re-draws of every randn call (normal_fill recomputes the last 16 values from fresh uniforms)
    std::vector<double> tails(ragged ? (size_t)n_iters * 2 * calls * 16 : 16);
    for (int it = 0; it < n_iters; ++it) {
        // main rollout: design calls silent (enable_noise=False), the action call rotates (acqfs.py:390-395)
        rotation_draw(g, N, d, u_main + (size_t)it * N, dims_main + (size_t)it * N * 2);
        for (int c = 0; c < calls; ++c)
            randn_f64_draw(g, N, nz_main + ((size_t)it * calls + c) * N,
                           ragged ? &tails[((size_t)(it * 2) * calls + c) * 16] : tails.data());
        // frozen "original" rollout: every policy call rotates (actor.py:585-593), then its cost calls
        for (int c = 0; c < calls; ++c)
            rotation_draw(g, N, d, u_orig + ((size_t)it * calls + c) * N, dims_orig + ((size_t)it * calls + c) * N * 2);
        for (int c = 0; c < calls; ++c)
            randn_f64_draw(g, N, nz_orig + ((size_t)it * calls + c) * N,
                           ragged ? &tails[((size_t)(it * 2 + 1) * calls + c) * 16] : tails.data());
        if (states_out) {
            uint8_t* dst = states_out + (size_t)it * kBlobBytes;
            memcpy(dst, state, kBlobBytes);
            g.store(dst);
        }
    }
    g.store(state);
    if (bulk) {
        // uniforms -> normals, one (iteration, rollout, call) block per task
        const long long n_blocks = (long long)n_iters * 2 * calls;
        auto work = [&](long long b0, long long b1) {
            for (long long b = b0; b < b1; ++b) {
                const long long it = b / (2 * calls), rem = b % (2 * calls);
                const int which = (int)(rem / calls), c = (int)(rem % calls);
                double* dst = (which ? nz_orig : nz_main) + ((size_t)it * calls + c) * N;
                randn_f64_finish(N, dst, ragged ? &tails[(size_t)b * 16] : nullptr);
            }
        };
        int nt = n_threads < 1 ? 1 : n_threads;
        if ((long long)nt > n_blocks) nt = (int)n_blocks;
        if (nt <= 1 || n_blocks * N < 65536) {
            work(0, n_blocks);
        } else {
            std::vector<std::thread> pool;
            for (int t = 0; t < nt; ++t)
                pool.emplace_back(work, n_blocks * t / nt, n_blocks * (t + 1) / nt);
            for (auto& th : pool) th.join();
        }
    }
    return 0;
}
