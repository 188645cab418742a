// The decode sampling step in ONE kernel: forbidden-token mask, presence / frequency penalties with decay, temperature,
// top-k, nucleus top-p, multinomial draw and the penalty-state update for the drawn token.
//
// Replaces, for one decode step, chirrup/worker.py:713-740 (forbidden tokens, `occurrence *= decay`,
// `logits -= alpha_presence + occurrence * frequency`), chirrup/utils/samplers.py:98-168 (`sample_logits_real_batch`: logits / T,
// softmax, ONE descending sort, keep rank < top_k and cumsum <= top_p with at least one token, renormalise, multinomial) and
// chirrup/worker.py:527-547 (`_batch_update_penalty`).  The reference runs a full sort of B x 65536 floats, two (B,V) fp32
// temporaries and B `.item()` syncs per step; here one CTA owns one row, reads the fp16 logits once and never sorts:
//
//   1. the penalised logits are fp16, so a row has at most 65536 distinct values: one pass builds a histogram of the row over
//      the order-preserving 16-bit keys of its values in shared memory (128 KiB, integer atomics: order-independent); the
//      probability mass of a key is a function of the key alone, evaluated as a 32-bit fixed-point integer, so every sum
//      below is an integer sum: deterministic whatever the reduction order;
//   2. the kept set of a descending sort is "every class above a boundary key, plus the first `take` elements of the boundary
//      class": one block-wide prefix scan of (count, count x mass) over the histogram in descending key order finds it;
//   3. the draw is u * (kept mass), located in the same scan as (key class, ordinal); a last pass over the row finds the
//      ordinal-th element with that key (ties are broken in thread-major order, which torch.sort leaves unspecified).
//
// mode 1 is the sampler the worker ACTUALLY imports (chirrup/worker.py:12: `sample_logits_rwkv_pip_compatible as
// sample_logits_real_batch`, chirrup/utils/samplers.py:171-255): softmax WITHOUT temperature; top-p keeps every token whose
// probability is >= the probability at the first sorted position where the cumulative sum reaches top_p (whole tie classes
// stay: a cutoff by VALUE); top-k keeps sort rank < k; THEN probs ** (1/T); multinomial.  In the histogram formulation: the
// boundary class is the first (descending) whose class mass brings the T = 1 cumulative mass to >= top_p * Z, it is kept
// whole unless top-k cuts into it, and the draw uses masses exp((l - max) / T) over the kept set (inv_T = fp16(1 / T), the
// reference's tensors are fp16).  mode 0 is the single-sort sampler of samplers.py:98-168 described above.
//
// Greedy rows (top_k == 1, top_p == 0 or temperature <= 0) take the arg-max with the lowest index, like b200_argmax_f16.
// The penalised logits are written back in place (fp32 math, one fp16 rounding per reference op), like the reference's decode_out.
#include "common.cuh"
#include "../../include/chirrup_b200.h"

namespace b200 {
namespace {

constexpr int ST = 1024;         // threads per row
constexpr int SCH = 8;           // 16-byte chunks per thread -> up to 1024 * 8 * 8 = 65536 columns
constexpr float FIX_ONE = 4294967296.0f;

typedef unsigned long long u64;

// Philox4x32-10 (Salmon et al. 2011): counter (row, step_lo, step_hi, 0), key = seed.
DEVI uint4 philox4x32(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; r++) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u;
        k.y += 0xBB67AE85u;
    }
    return c;
}

// fp16 bit pattern -> 16-bit key with the same order as the value (larger value = larger key); key 0 is reserved for padding
// (it would be a negative NaN payload; every finite value and -inf map to keys >= 0x03FF).
DEVI uint32_t key_of(uint32_t h) { return (h & 0x8000u) ? (~h & 0xFFFFu) : (h | 0x8000u); }
DEVI float val_of_key(uint32_t k) {
    const uint32_t h = (k & 0x8000u) ? (k & 0x7FFFu) : (~k & 0xFFFFu);
    return __half2float(__ushort_as_half(static_cast<unsigned short>(h)));
}
struct Scratch {
    u64 w64[ST / 32];
    uint32_t w32[ST / 32];
    u64 c_star, m_star, mass_prime, ord;
    uint32_t kprime, c_prime, kd, b32;
};

// Block-wide sums (every thread gets the total).  Integer sums: order-independent, hence deterministic.
DEVI void block_sum_cm(uint32_t& cnt, u64& mass, Scratch& s) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        mass += __shfl_xor_sync(0xffffffffu, mass, o);
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();  // previous readers of the scratch are done
    if (lane == 0) s.w32[warp] = cnt, s.w64[warp] = mass;
    __syncthreads();
    uint32_t c = lane < ST / 32 ? s.w32[lane] : 0u;
    u64 m = lane < ST / 32 ? s.w64[lane] : 0ull;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        c += __shfl_xor_sync(0xffffffffu, c, o);
        m += __shfl_xor_sync(0xffffffffu, m, o);
    }
    cnt = c;
    mass = m;
}

// Exclusive prefix of one u64 per thread in thread order; `total` = block sum.
DEVI u64 block_excl_scan(u64 v, u64& total, Scratch& s) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    u64 inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const u64 t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads();
    if (lane == 31) s.w64[warp] = inc;
    __syncthreads();
    u64 wsum = lane < ST / 32 ? s.w64[lane] : 0ull;
    u64 winc = wsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const u64 t = __shfl_up_sync(0xffffffffu, winc, o);
        if (lane >= o) winc += t;
    }
    total = __shfl_sync(0xffffffffu, winc, ST / 32 - 1);
    const u64 wexc = __shfl_sync(0xffffffffu, winc - wsum, warp);
    return wexc + inc - v;
}

DEVI u64 block_max_u64(u64 v, Scratch& s) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const u64 t = __shfl_xor_sync(0xffffffffu, v, o);
        v = t > v ? t : v;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) s.w64[warp] = v;
    __syncthreads();
    u64 m = lane < ST / 32 ? s.w64[lane] : 0ull;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const u64 t = __shfl_xor_sync(0xffffffffu, m, o);
        m = t > m ? t : m;
    }
    return m;
}

struct SampleArgs {
    int rows, V;
    __half* logits;
    long ld;
    const int* slot;
    float* occurrence;
    float* presence;
    const float *decay, *freq, *pres_pen, *temp, *top_p;
    const int* top_k;
    const int* forbidden;
    int max_forbidden;
    const int* no_penalty;
    int n_no_penalty;
    u64 seed, step;
    int* kept;
    int* out;
    int mode;   // 0: samplers.py:98-168 (temperature first, nucleus by cumulative mass); 1: samplers.py:171-255 (rwkv pip)
};

constexpr int NKEY = 65536;          // distinct fp16 bit patterns
constexpr int HWORDS = NKEY / 2;     // the per-key count histogram: two 16-bit counters per 32-bit word (128 KiB)
constexpr int WPT = HWORDS / ST;     // histogram words owned by one thread in the scan (64 words = 128 keys)

DEVI uint32_t hist_count(const uint32_t* hist, uint32_t key) { return (hist[key >> 1] >> (16 * (key & 1u))) & 0xFFFFu; }

__global__ void __launch_bounds__(ST, 1) sample_rows_kernel(const SampleArgs a) {
    __shared__ Scratch s;
    extern __shared__ __align__(16) uint32_t hist[];
    pdl_launch_dependents();
    pdl_wait();
    const int row = blockIdx.x, t = threadIdx.x;
    const int V = a.V;
    const int sl = a.slot ? a.slot[row] : row;
    __half* lrow = a.logits + (size_t)row * a.ld;
    const float temp = a.temp ? a.temp[sl] : 1.f;
    const float top_p = a.top_p ? a.top_p[sl] : 1.f;
    const int top_k = a.top_k ? a.top_k[sl] : 0;
    const bool greedy = top_k == 1 || !(top_p > 0.f) || !(temp > 0.f);

    // ---- forbidden tokens: decode_out[..][tok] -= 1e10 on the fp16 row (worker.py:719-722)
    if (a.forbidden != nullptr && t < a.max_forbidden) {
        const int tok = a.forbidden[(size_t)row * a.max_forbidden + t];
        if (tok >= 0 && tok < V) lrow[tok] = __float2half_rn(__half2float(lrow[tok]) - 1e10f);
    }
    if (!greedy) {
        for (int i = t; i < HWORDS / 4; i += ST) reinterpret_cast<uint4*>(hist)[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncthreads();

    // ---- one pass over the row: decay + penalties (worker.py:724-728) written back in place, arg-max, per-key histogram
    const bool pen = a.occurrence != nullptr;
    const float decay = pen ? a.decay[sl] : 1.f, freq = pen ? a.freq[sl] : 0.f;
    u64 best = 0ull;  // (key << 32) | ~index : max = highest value, lowest index
    uint32_t kmin = 0xFFFFu;
#pragma unroll 1
    for (int j = 0; j < SCH; j++) {
        const int c = (j * ST + t) * 8;
        if (c >= V) break;
        int4 q = *reinterpret_cast<const int4*>(lrow + c);
        __half* h = reinterpret_cast<__half*>(&q);
        if (pen) {
            float* op = a.occurrence + (size_t)sl * V + c;
            const float* pp = a.presence + (size_t)sl * V + c;
            const float4 o0 = *reinterpret_cast<const float4*>(op), o1 = *reinterpret_cast<const float4*>(op + 4);
            const float4 p0 = *reinterpret_cast<const float4*>(pp), p1 = *reinterpret_cast<const float4*>(pp + 4);
            float o[8] = {o0.x, o0.y, o0.z, o0.w, o1.x, o1.y, o1.z, o1.w};
            const float al[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
#pragma unroll
            for (int e = 0; e < 8; e++) {
                o[e] = __fmul_rn(o[e], decay);
                const float pv = __fadd_rn(al[e], __fmul_rn(o[e], freq));
                h[e] = __float2half_rn(__fsub_rn(__half2float(h[e]), pv));
            }
            *reinterpret_cast<float4*>(op) = make_float4(o[0], o[1], o[2], o[3]);
            *reinterpret_cast<float4*>(op + 4) = make_float4(o[4], o[5], o[6], o[7]);
            *reinterpret_cast<int4*>(lrow + c) = q;
        }
        const unsigned short* hb = reinterpret_cast<const unsigned short*>(&q);
#pragma unroll
        for (int e = 0; e < 8; e++) {
            const uint32_t k = key_of(hb[e]);
            const u64 bk = ((u64)k << 32) | (u64)(0xFFFFFFFFu - (uint32_t)(c + e));
            best = bk > best ? bk : best;
            kmin = k < kmin ? k : kmin;
            if (!greedy) atomicAdd(&hist[k >> 1], 1u << (16 * (k & 1u)));  // integer: order-independent
        }
    }
    best = block_max_u64(best, s);
    const uint32_t kmax = (uint32_t)(best >> 32);
    int tok = (int)(0xFFFFFFFFu - (uint32_t)(best & 0xFFFFFFFFull));
    int n_kept = 1;

    if (!greedy) {
        const uint32_t kmin_all = 0xFFFFu - (uint32_t)block_max_u64((u64)(0xFFFFu - kmin), s);
        const bool pip = a.mode == 1;
        const float inv_t = pip ? __half2float(__float2half_rn(1.0f / temp)) : 1.0f / temp;   // :244 on fp16 tensors
        const float vmax = val_of_key(kmax);
        const float xm = vmax * inv_t;
        // unnormalised probability exp(l/T - max/T) of a key as 32-bit fixed point (1.0 -> 2^32 - 1, exp(-inf) -> 0): what the
        // draw uses in both modes, and what the filters use in mode 0
        auto mass_of_key = [&](uint32_t key) -> u64 {
            const float e = exp2f((val_of_key(key) * inv_t - xm) * 1.4426950408889634f);
            return (u64)__float2uint_rd(e * FIX_ONE);
        };
        // mode 1 filters on the UNtempered distribution softmax(logits) (samplers.py:198)
        const bool two_mass = pip && inv_t != 1.0f;
        auto fmass_of_key = [&](uint32_t key) -> u64 {
            if (!two_mass) return mass_of_key(key);
            const float e = exp2f((val_of_key(key) - vmax) * 1.4426950408889634f);
            return (u64)__float2uint_rd(e * FIX_ONE);
        };
        const uint4 rnd = philox4x32(make_uint4((uint32_t)row, (uint32_t)a.step, (uint32_t)(a.step >> 32), 0u),
                                     make_uint2((uint32_t)a.seed, (uint32_t)(a.seed >> 32)));
        const u64 r64 = ((u64)rnd.x << 32) | rnd.y;
        uint32_t k_d;  // the drawn key class ...
        u64 ord;       // ... and which of its elements (thread-major order)
        if (kmin_all == kmax) {
            // every logit equal (the one case in which a 16-bit counter can wrap): uniform over the first `take` elements
            const u64 m1 = mass_of_key(kmax);
            const u64 k_eff = top_k > 0 ? (u64)min(top_k, V) : (u64)V;
            u64 take = top_p < 1.0f ? (u64)((double)V * (double)top_p) : (u64)V;  // cumsum_j = (j+1)/V <= top_p
            if (pip) take = (u64)V;                                               // one tie class: the cutoff value keeps it whole
            take = take < k_eff ? take : k_eff;
            take = take ? take : 1ull;
            (void)m1;
            k_d = kmax;
            ord = __umul64hi(r64, take);
            n_kept = (int)take;
        } else {
            // ---- histogram scan in DESCENDING key order: thread t owns words w_hi, w_hi-1, .., w_hi-63 (keys 2w+1, 2w)
            const int lane = t & 31;
            const uint32_t w_hi = HWORDS - 1 - t * WPT;
            u64 cnt = 0ull, mass = 0ull, dmass = 0ull;   // mass: what the filters see; dmass: what the draw sees (mode 1, T != 1)
#pragma unroll 4
            for (int i = 0; i < WPT; i++) {
                const uint32_t w = w_hi - ((i + lane) & (WPT - 1));  // rotated per lane: conflict-free; sums do not care
                const uint32_t word = hist[w];
                if (word) {
                    const uint32_t c1 = word >> 16, c0 = word & 0xFFFFu;
                    cnt += c1 + c0;
                    if (c1) mass += c1 * fmass_of_key(2 * w + 1);
                    if (c0) mass += c0 * fmass_of_key(2 * w);
                    if (two_mass) {
                        if (c1) dmass += c1 * mass_of_key(2 * w + 1);
                        if (c0) dmass += c0 * mass_of_key(2 * w);
                    }
                }
            }
            if (t == 0) {
                s.kprime = 0u;
                s.c_prime = 0u;
                s.mass_prime = 0ull;
            }
            u64 n_valid, Z, Zd;
            const u64 pc = block_excl_scan(cnt, n_valid, s);
            const u64 pm = block_excl_scan(mass, Z, s);
            const u64 pd = two_mass ? block_excl_scan(dmass, Zd, s) : pm;   // draw-mass prefix
            if (!two_mass) dmass = mass, Zd = Z;
            const u64 k_eff = top_k > 0 ? min((u64)top_k, n_valid) : n_valid;
            u64 P = top_p < 1.0f ? (u64)((double)Z * (double)top_p) : ~0ull;
            if (pip && P == 0ull) P = 1ull;
            if (t == 0) s.c_star = n_valid, s.m_star = Zd;
            __syncthreads();
            // the first key (descending) at which a filter bites.  mode 0: the class no longer fits whole — rank >= k_eff or
            // cumulative mass > P.  mode 1: rank >= k_eff, or the class brings the cumulative mass to >= P (it is then the LAST
            // class kept, and kept whole: cutoff by value, samplers.py:209-216)
            const bool before = pip ? (pc > k_eff || pm >= P) : (pc > k_eff || pm > P);
            const bool within = pip ? (pc + cnt > k_eff || pm + mass >= P) : (pc + cnt > k_eff || pm + mass > P);
            if (!before && within) {
                u64 cc = pc, cm = pm, cd = pd;
                for (int i = 0; i < 2 * WPT; i++) {
                    const uint32_t key = 2 * w_hi + 1 - i;
                    const u64 c = hist_count(hist, key);
                    if (!c) continue;
                    const u64 m1 = fmass_of_key(key);
                    if (cc + c > k_eff || (pip ? cm + c * m1 >= P : cm + c * m1 > P)) {
                        s.kprime = key;
                        s.c_prime = (uint32_t)c;
                        s.mass_prime = pip ? mass_of_key(key) : m1;     // per-element DRAW mass of the boundary class
                        s.c_star = cc;
                        s.m_star = pip ? cd : cm;                        // draw mass of the classes kept whole above it
                        s.b32 = (pip && !(cc + c > k_eff)) ? 1u : 0u;    // mode 1, nucleus boundary: the whole class stays
                        break;
                    }
                    cc += c;
                    cm += c * m1;
                    if (two_mass) cd += c * mass_of_key(key); else cd = cm;
                }
            }
            __syncthreads();
            const uint32_t kprime = s.kprime;
            const u64 c_star = s.c_star, m_star = s.m_star, mass_prime = s.mass_prime;
            u64 take = 0ull;
            if (kprime) {  // partially kept class: equal masses, so `take` of its elements
                take = min((u64)s.c_prime, k_eff - c_star);
                if (pip) {
                    if (s.b32) take = (u64)s.c_prime;
                } else {
                    const u64 rem_p = mass_prime ? (P - m_star) / mass_prime : (u64)s.c_prime;
                    take = take < rem_p ? take : rem_p;
                }
                if (c_star == 0ull && take == 0ull) take = 1ull;  // min_tokens_to_keep = 1 (samplers.py:105, :145-146)
            }
            const u64 m_kept = m_star + take * mass_prime;
            n_kept = (int)(c_star + take);
            // ---- the draw: target in [0, m_kept) -> key class + ordinal inside the class
            const u64 target = __umul64hi(r64, m_kept);
            if (target >= m_star) {
                if (t == 0) s.kd = kprime, s.ord = mass_prime ? (target - m_star) / mass_prime : 0ull;
            } else if (pd <= target && target < pd + dmass) {
                u64 cm = pd;
                for (int i = 0; i < 2 * WPT; i++) {
                    const uint32_t key = 2 * w_hi + 1 - i;
                    const u64 c = hist_count(hist, key);
                    if (!c) continue;
                    const u64 m1 = mass_of_key(key);
                    if (target < cm + c * m1) {
                        s.kd = key;
                        s.ord = (target - cm) / m1;
                        break;
                    }
                    cm += c * m1;
                }
            }
            __syncthreads();
            k_d = s.kd;
            ord = s.ord;
        }
        // ---- the ord-th element (thread-major order) carrying the drawn key: one more pass over the (penalised) row
        u64 mine = 0ull, tot;
#pragma unroll 1
        for (int j = 0; j < SCH; j++) {
            const int c = (j * ST + t) * 8;
            if (c >= V) break;
            const int4 q = *reinterpret_cast<const int4*>(lrow + c);
            const unsigned short* hb = reinterpret_cast<const unsigned short*>(&q);
#pragma unroll
            for (int e = 0; e < 8; e++) mine += key_of(hb[e]) == k_d ? 1ull : 0ull;
        }
        if (t == 0) s.b32 = (uint32_t)tok;
        const u64 before = block_excl_scan(mine, tot, s);
        if (before <= ord && ord < before + mine) {
            u64 seen = before;
#pragma unroll 1
            for (int j = 0; j < SCH; j++) {
                const int c = (j * ST + t) * 8;
                if (c >= V) break;
                const int4 q = *reinterpret_cast<const int4*>(lrow + c);
                const unsigned short* hb = reinterpret_cast<const unsigned short*>(&q);
#pragma unroll
                for (int e = 0; e < 8; e++) {
                    if (key_of(hb[e]) == k_d) {
                        if (seen == ord) s.b32 = (uint32_t)(c + e);
                        seen++;
                    }
                }
            }
        }
        __syncthreads();
        tok = (int)s.b32;
    }

    // ---- result + penalty state of the drawn token (_batch_update_penalty, worker.py:527-547; weight 0 for no-penalty ids)
    if (t == 0) {
        a.out[row] = tok;
        if (a.kept) a.kept[row] = n_kept;
        if (pen) {
            float w = 1.0f;
            for (int i = 0; i < a.n_no_penalty; i++)
                if (a.no_penalty[i] == tok) w = 0.0f;
            a.occurrence[(size_t)sl * V + tok] += w;
            a.presence[(size_t)sl * V + tok] = a.pres_pen[sl];
        }
    }
}

}  // namespace
}  // namespace b200

extern "C" int b200_sample_f16(int rows, int V, void* logits, long ld, const b200_sample_params* p, int* out, void* stream) {
    using namespace b200;
    if (rows < 0 || V <= 0 || (V & 7) || (ld & 7) || !p || !out || !logits) return B200_ERR_BAD_ARG;
    if (V > ST * SCH * 8) return B200_ERR_UNSUPPORTED;
    if ((p->occurrence == nullptr) != (p->presence == nullptr)) return B200_ERR_BAD_ARG;
    if (p->occurrence && (!p->penalty_decay || !p->frequency_penalty || !p->presence_penalty)) return B200_ERR_BAD_ARG;
    if (p->max_forbidden < 0 || p->max_forbidden > ST || p->n_no_penalty < 0 || p->mode < 0 || p->mode > 1) return B200_ERR_BAD_ARG;
    if ((reinterpret_cast<uintptr_t>(logits) | reinterpret_cast<uintptr_t>(p->occurrence) | reinterpret_cast<uintptr_t>(p->presence)) & 15)
        return B200_ERR_BAD_ARG;
    if (rows == 0) return 0;
    SampleArgs a;
    a.rows = rows; a.V = V; a.logits = static_cast<__half*>(logits); a.ld = ld;
    a.slot = p->slot; a.occurrence = p->occurrence; a.presence = p->presence;
    a.decay = p->penalty_decay; a.freq = p->frequency_penalty; a.pres_pen = p->presence_penalty;
    a.temp = p->temperature; a.top_p = p->top_p; a.top_k = p->top_k;
    a.forbidden = p->max_forbidden > 0 ? p->forbidden : nullptr; a.max_forbidden = p->max_forbidden;
    a.no_penalty = p->no_penalty; a.n_no_penalty = p->no_penalty ? p->n_no_penalty : 0;
    a.seed = p->seed; a.step = p->step; a.kept = p->kept; a.out = out; a.mode = p->mode;
    const size_t smem = (size_t)HWORDS * 4;  // the per-key histogram
    cudaError_t e = cudaFuncSetAttribute(sample_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    return launch_kernel(sample_rows_kernel, dim3(rows), dim3(ST), smem, static_cast<cudaStream_t>(stream), false, 0, a);
}
