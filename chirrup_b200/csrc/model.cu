// The batched RWKV-7 step as one sequence of sm_100a kernel launches on a caller-supplied stream.
// Follows RWKV_x070.forward_seq_batch_seperate (/root/reference/Albatross/rwkv7.py:555-563 -> _pre :503-509,
// _layers :511-538, _post :540-553) with TMix = rwkv7.py:618-649 and CMix = :673-679.
#include <cstdio>
#include <cstdlib>
#include <new>

#include "common.cuh"
#include "glue.h"
#include "../../include/chirrup_b200.h"

struct b200_model {
    int L, C, V, H, F, d_w, d_a, d_v, d_g;
    const __half* gw[4];
    const __half** lw;  // L * B200_LW_COUNT
    const __half** ffn_vt;  // optional: L pointers to ffn.value.weight in the reference's (F, C) layout (bsz-1 sparse product)
    int last_launches;
};

namespace b200 {
namespace {

inline long align_up(long x, long a) { return (x + a - 1) / a * a; }

// Workspace carve-up for M tokens.
struct Ws {
    int* flags;                  // 16 ints for chained launches (gemm_pair.cu), zero between launches
    int* counters_and_partials;  // linear workspace
    long lin_bytes;
    __half *x, *mix[6], *r, *k, *v, *w, *a, *g, *vg, *k2, *v2, *an, *bn, *y, *yg, *vfirst, *kx, *lora[4], *hidden, *xo;
    long total;
};

Ws carve(const b200_model* m, long M, void* base) {
    Ws ws;
    uint8_t* p = static_cast<uint8_t*>(base);
    long off = 0;
    auto take = [&](long bytes) {
        uint8_t* q = p ? p + off : nullptr;
        off += align_up(bytes, 256);
        return q;
    };
    ws.flags = reinterpret_cast<int*>(take(256));  // first, so that its place does not depend on M (chained launches keep it zero)
    ws.lin_bytes = b200_linear_workspace_bytes((int)M, m->V);
    ws.counters_and_partials = reinterpret_cast<int*>(take(ws.lin_bytes));
    const long mc = M * m->C * 2;
    __half** bufs[] = {&ws.x,  &ws.mix[0], &ws.mix[1], &ws.mix[2], &ws.mix[3], &ws.mix[4], &ws.mix[5], &ws.r, &ws.k,
                       &ws.v,  &ws.w,      &ws.a,      &ws.g,      &ws.vg,     &ws.k2,     &ws.v2,     &ws.an, &ws.bn,
                       &ws.y,  &ws.yg,     &ws.vfirst, &ws.kx,     &ws.xo};
    for (auto b : bufs) *b = reinterpret_cast<__half*>(take(mc));
    int dmax = m->d_w;
    if (m->d_a > dmax) dmax = m->d_a;
    if (m->d_v > dmax) dmax = m->d_v;
    if (m->d_g > dmax) dmax = m->d_g;
    for (int i = 0; i < 4; i++) ws.lora[i] = reinterpret_cast<__half*>(take(M * align_up(dmax, 8) * 2));
    ws.hidden = reinterpret_cast<__half*>(take(M * (long)m->F * 2));
    ws.total = off;
    return ws;
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" int b200_model_create(int n_layer, int n_embd, int vocab_size, int d_w, int d_a, int d_v, int d_g,
                                 const void* const* global_w, const void* const* layer_w, b200_model** out) {
    if (!out || !global_w || !layer_w || n_layer <= 0 || n_embd <= 0 || (n_embd % 64) || vocab_size <= 0 || (vocab_size & 7))
        return B200_ERR_BAD_ARG;
    if (n_embd > 4096) return B200_ERR_UNSUPPORTED;  // ln_shift_mix: one thread per 8 channels, 512 threads (RWKV-7 g1: C <= 4096)
    if (d_w <= 0 || d_a <= 0 || d_v <= 0 || d_g <= 0 || (d_w & 7) || (d_a & 7) || (d_v & 7) || (d_g & 7)) return B200_ERR_BAD_ARG;
    b200_model* m = new (std::nothrow) b200_model();
    if (!m) return B200_ERR_BAD_ARG;
    m->L = n_layer; m->C = n_embd; m->V = vocab_size; m->H = n_embd / 64; m->F = 4 * n_embd;
    m->d_w = d_w; m->d_a = d_a; m->d_v = d_v; m->d_g = d_g;
    for (int i = 0; i < 4; i++) m->gw[i] = static_cast<const __half*>(global_w[i]);
    m->lw = new (std::nothrow) const __half*[(size_t)n_layer * B200_LW_COUNT];
    for (long i = 0; i < (long)n_layer * B200_LW_COUNT; i++) m->lw[i] = static_cast<const __half*>(layer_w[i]);
    m->ffn_vt = nullptr;
    m->last_launches = 0;
    *out = m;
    return 0;
}

extern "C" int b200_model_set_ffn_value_rows(b200_model* m, const void* const* ffn_value_fc) {
    if (!m) return B200_ERR_BAD_ARG;
    delete[] m->ffn_vt;
    m->ffn_vt = nullptr;
    if (!ffn_value_fc) return 0;
    m->ffn_vt = new (std::nothrow) const __half*[(size_t)m->L];
    if (!m->ffn_vt) return B200_ERR_BAD_ARG;
    for (int i = 0; i < m->L; i++) m->ffn_vt[i] = static_cast<const __half*>(ffn_value_fc[i]);
    return 0;
}

extern "C" void b200_model_destroy(b200_model* m) {
    if (!m) return;
    delete[] m->lw;
    delete[] m->ffn_vt;
    delete m;
}

extern "C" long b200_model_workspace_bytes(const b200_model* m, int max_tokens) {
    if (!m || max_tokens < 0) return B200_ERR_BAD_ARG;
    return carve(m, max_tokens < 1 ? 1 : max_tokens, nullptr).total;
}

extern "C" int b200_model_last_launch_count(const b200_model* m) { return m ? m->last_launches : 0; }

extern "C" int b200_model_forward(b200_model* m, int B, int T, const int* tokens, void* state0, long s0_layer_stride,
                                  long s0_which_stride, void* state1, long s1_layer_stride, int n_slots, int* elapsed,
                                  const int* slot_idx, void* logits, int full_output, void* workspace, long workspace_bytes,
                                  void* stream_) {
    if (!m || B < 0 || T < 0) return B200_ERR_BAD_ARG;
    if (B == 0 || T == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream_);
    const int C = m->C, H = m->H, F = m->F, L = m->L;
    const long M = (long)B * T;
    Ws ws = carve(m, M, workspace);
    if (!workspace || workspace_bytes < ws.total) return B200_ERR_WORKSPACE;
    if (!slot_idx) n_slots = B;
    int n = 0, rc = 0;
    static const bool debug_sync = getenv("B200_DEBUG_SYNC") != nullptr;
    const bool unfused_tmix = getenv("B200_UNFUSED_TMIX") != nullptr;  // A/B switch: the three-kernel chain of glue.cu
    // Programmatic dependent launch for every kernel after the first (the first one orders the step after whatever the caller
    // enqueued before it); every kernel of the step calls griddepcontrol.wait before touching its predecessors' outputs.
    const bool pdl = getenv("B200_NO_PDL") == nullptr;
    const bool wkv_sequential = getenv("B200_WKV_SEQUENTIAL") != nullptr;
    const bool no_spmv = getenv("B200_NO_SPMV") != nullptr;  // A/B switch: dense FFN value product at B = T = 1
    const bool no_group = getenv("B200_NO_PAIR") != nullptr;  // A/B switch: the one-CTA-per-tile GEMM path of gemm.cu
#define RUN(call)                                                                                     \
    do {                                                                                              \
        rc = (call);                                                                                  \
        n++;                                                                                          \
        if (!rc && debug_sync) rc = (int)cudaStreamSynchronize(st);                                   \
        if (rc) {                                                                                     \
            if (debug_sync) fprintf(stderr, "[b200] launch #%d failed rc=%d: %s\n", n, rc, #call);    \
            return rc;                                                                                \
        }                                                                                             \
    } while (0)
    // Timing aid (results are garbage): B200_DEBUG_SKIP is a bit mask of launches to leave out, so that the in-graph marginal cost
    // of each kernel of the layer can be read off the step time (tools/knockout.sh): 1 ln_shift_mix6, 2 r/k/v + LoRA-down group,
    // 4 LoRA-up group, 8 fused WKV, 16 output projection, 32 ln_shift_mix1, 64 FFN key, 128 FFN value, 256 ln_out + head.
    const int skip = getenv("B200_DEBUG_SKIP") ? atoi(getenv("B200_DEBUG_SKIP")) : 0;
#define RUN_T(tag, call)            \
    do {                            \
        if (!(skip & (tag))) RUN(call); \
    } while (0)
    auto lin = [&](long rows, int N, int K, const __half* A, const __half* W, __half* out, const __half* bias,
                   const __half* res, int epi, int* defer_splits = nullptr) {
        return linear_f16_launch((int)rows, N, K, A, K, W, K, out, N, bias, res, epi, ws.counters_and_partials, ws.lin_bytes, st,
                                 pdl, defer_splits);
    };
    // Decode (T == 1): the two residual-add GEMMs (output projection, FFN value) may leave their K split unreduced in the
    // linear workspace; the LayerNorm/shift/mix kernel that consumes x next finishes the add (glue.h).  The workspace is free
    // in between: the consumer is always the very next launch.
    const bool may_defer = T == 1 && getenv("B200_NO_DEFER") == nullptr;
    const float* partials = reinterpret_cast<const float*>(ws.counters_and_partials);
    int def_o = 0, def_f = 0;  // pending splits of the output projection / of the previous layer's FFN value
    __half* s0 = static_cast<__half*>(state0);
    __half* s1 = static_cast<__half*>(state1);

    RUN(launch_embed_gather((int)M, C, m->V, tokens, m->gw[0], ws.x, st, /*pdl=*/false));
    for (int l = 0; l < L; l++) {
        const __half** w = m->lw + (size_t)l * B200_LW_COUNT;
        // ---- TMix
        MixPtrs<6> mp;
        for (int i = 0; i < 6; i++) {
            mp.mix[i] = w[B200_LW_X_R + i];
            mp.out[i] = ws.mix[i];
        }
        __half *xr = ws.mix[0], *xw = ws.mix[1], *xk = ws.mix[2], *xv = ws.mix[3], *xa = ws.mix[4], *xg = ws.mix[5];
        RUN_T(1, launch_ln_shift_mix6(B, T, C, ws.x, w[B200_LW_LN1_W], w[B200_LW_LN1_B], s0 + l * s0_layer_stride, C, slot_idx, mp, st, pdl,
                                 def_f ? partials : nullptr, def_f));
        def_f = 0;
        __half* vbuf = (l == 0) ? ws.vfirst : ws.v;  // layer 0: v_first = v (rwkv7.py:636)
        // r, k, v and the four LoRA down-projections read only the six mixes: ONE grouped launch (7 problems; 6 on layer 0, which
        // has no v-gate, rwkv7.py:636), then the four up-projections in another
        // (w: tanh, +w0 ; a: +a0, sigmoid ; g: sigmoid inside ; v-gate: +v0, sigmoid — rwkv7.py:625-630, :637)
        {
            const int nl = (l > 0) ? 4 : 3;
            LinearDesc dn[7] = {
                {C, C, xr, C, w[B200_LW_RECEPTANCE], C, ws.r, C, nullptr, nullptr, B200_EPI_NONE},
                {C, C, xk, C, w[B200_LW_KEY], C, ws.k, C, nullptr, nullptr, B200_EPI_NONE},
                {C, C, xv, C, w[B200_LW_VALUE], C, vbuf, C, nullptr, nullptr, B200_EPI_NONE},
                {m->d_w, C, xw, C, w[B200_LW_W1], C, ws.lora[0], m->d_w, nullptr, nullptr, B200_EPI_TANH},
                {m->d_a, C, xa, C, w[B200_LW_A1], C, ws.lora[1], m->d_a, nullptr, nullptr, B200_EPI_NONE},
                {m->d_g, C, xg, C, w[B200_LW_G1], C, ws.lora[2], m->d_g, nullptr, nullptr, B200_EPI_SIGMOID},
                {m->d_v, C, xv, C, w[B200_LW_V1], C, ws.lora[3], m->d_v, nullptr, nullptr, B200_EPI_NONE}};
            if (M > 128 && !no_group) {
                RUN_T(2, linear_group_launch(3 + nl, dn, (int)M, ws.counters_and_partials, ws.lin_bytes, st, pdl));
            } else {
                for (int i = 0; i < 3; i++)
                    RUN(linear_f16_launch((int)M, dn[i].N, dn[i].K, dn[i].A, dn[i].lda, dn[i].W, dn[i].ldw, dn[i].out, dn[i].ldo,
                                          nullptr, nullptr, dn[i].epi, ws.counters_and_partials, ws.lin_bytes, st, pdl));
                RUN(linear_group_launch(nl, dn + 3, (int)M, ws.counters_and_partials, ws.lin_bytes, st, pdl));
            }
            LinearDesc up[4] = {
                {C, m->d_w, ws.lora[0], m->d_w, w[B200_LW_W2], m->d_w, ws.w, C, w[B200_LW_W0], nullptr, B200_EPI_NONE},
                {C, m->d_a, ws.lora[1], m->d_a, w[B200_LW_A2], m->d_a, ws.a, C, w[B200_LW_A0], nullptr, B200_EPI_SIGMOID},
                {C, m->d_g, ws.lora[2], m->d_g, w[B200_LW_G2], m->d_g, ws.g, C, nullptr, nullptr, B200_EPI_NONE},
                {C, m->d_v, ws.lora[3], m->d_v, w[B200_LW_V2], m->d_v, ws.vg, C, w[B200_LW_V0], nullptr, B200_EPI_SIGMOID}};
            RUN_T(4, linear_group_launch(nl, up, (int)M, ws.counters_and_partials, ws.lin_bytes, st, pdl));
        }
        if (!unfused_tmix) {
            // prep + WKV-7 + GroupNorm/bonus/gate in one kernel per (sequence, head) tile
            WkvFusedArgs fa;
            fa.B = B; fa.T = T; fa.C = C; fa.H = H; fa.rows_per_slot = 0;
            fa.r = ws.r; fa.w = ws.w; fa.k = ws.k; fa.v = vbuf; fa.a = ws.a; fa.g = ws.g;
            fa.vgate = l > 0 ? ws.vg : nullptr;
            fa.v_first = l > 0 ? ws.vfirst : nullptr;
            fa.k_k = w[B200_LW_K_K]; fa.k_a = w[B200_LW_K_A]; fa.r_k = w[B200_LW_R_K];
            fa.ln_w = w[B200_LW_LNX_W]; fa.ln_b = w[B200_LW_LNX_B];
            fa.out = ws.yg; fa.elapsed_t = elapsed; fa.slot_idx = slot_idx;
            // prefill (T >= 32): the chunked tensor-core form of the recurrence; decode and short T: the sequential kernels,
            // bit-identical to the reference kernel (B200_WKV_SEQUENTIAL=1 forces them everywhere)
            if (T >= 32 && !wkv_sequential)
                RUN_T(8, launch_wkv7_chunk(fa, s1 + l * s1_layer_stride, (long)C * 64, n_slots, st, pdl));
            else
                RUN_T(8, launch_wkv7_fused(fa, s1 + l * s1_layer_stride, (long)C * 64, n_slots, st, pdl));
        } else {
            TmixPrepArgs pa;
            pa.M = (int)M; pa.C = C;
            pa.k = ws.k; pa.v = vbuf; pa.a = ws.a;
            pa.vgate = l > 0 ? ws.vg : nullptr;
            pa.v_first = l > 0 ? ws.vfirst : nullptr;
            pa.k_k = w[B200_LW_K_K]; pa.k_a = w[B200_LW_K_A];
            pa.k_out = ws.k2; pa.v_out = ws.v2; pa.a_out = ws.an; pa.b_out = ws.bn;
            RUN(launch_tmix_prep(pa, st, pdl));
            const __half* vv = l > 0 ? ws.v2 : ws.vfirst;
            RUN(launch_wkv7_seq(B, T, C, H, s1 + l * s1_layer_stride, (long)C * 64, n_slots, ws.r, ws.w, ws.k2, vv, ws.an, ws.bn,
                                ws.y, elapsed, slot_idx, st, pdl));
            TmixPostArgs po;
            po.M = (int)M; po.C = C;
            po.y = ws.y; po.r = ws.r; po.k = ws.k2; po.v = vv; po.g = ws.g;
            po.ln_w = w[B200_LW_LNX_W]; po.ln_b = w[B200_LW_LNX_B]; po.r_k = w[B200_LW_R_K];
            po.out = ws.yg;
            RUN(launch_tmix_post(po, st, pdl));
        }
        RUN_T(16, lin(M, C, C, ws.yg, w[B200_LW_OUTPUT], ws.x, nullptr, ws.x, B200_EPI_RESADD, may_defer ? &def_o : nullptr));
        // ---- CMix
        MixPtrs<1> mc;
        mc.mix[0] = w[B200_LW_FFN_X_K];
        mc.out[0] = ws.kx;
        RUN_T(32, launch_ln_shift_mix1(B, T, C, ws.x, w[B200_LW_LN2_W], w[B200_LW_LN2_B], s0 + l * s0_layer_stride + s0_which_stride, C,
                                 slot_idx, mc, st, pdl, def_o ? partials : nullptr, def_o));
        def_o = 0;
        bool chained = false;
        if (may_defer && l + 1 < L && M > 128 && M <= 256 && !no_group && !skip) {
            // FFN key + FFN value in one persistent launch (value partials deferred to the next layer's ln_shift_mix)
            rc = ffn_chain_launch((int)M, C, F, ws.kx, w[B200_LW_FFN_KEY], ws.hidden, w[B200_LW_FFN_VALUE],
                                  reinterpret_cast<float*>(ws.counters_and_partials), ws.lin_bytes, ws.flags, &def_f, st, pdl);
            if (rc == 0) {
                chained = true;
                n++;
                if (debug_sync) rc = (int)cudaStreamSynchronize(st);
            }
            if (rc != 0 && rc != B200_ERR_UNSUPPORTED) return rc;
        }
        if (M == 1 && m->ffn_vt && !no_spmv) {
            // bsz-1 path (RWKV_x070_CMix_one, rwkv7.py:653-662): the ReLU^2 output is sparse, so the value product reads only
            // the rows of the (F, C) matrix whose key activation is non-zero — rwkv_mm_sparsity / SPMV_OP, with the residual add
            RUN_T(64, lin(M, F, C, ws.kx, w[B200_LW_FFN_KEY], ws.hidden, nullptr, nullptr, B200_EPI_RELU_SQ));
            RUN_T(128, launch_spmv(F, C, ws.hidden, m->ffn_vt[l], ws.x, ws.x, st, pdl));
        } else if (!chained) {
            RUN_T(64, lin(M, F, C, ws.kx, w[B200_LW_FFN_KEY], ws.hidden, nullptr, nullptr, B200_EPI_RELU_SQ));
            // the last layer's x goes to ln_out, which takes finished rows
            RUN_T(128, lin(M, C, F, ws.hidden, w[B200_LW_FFN_VALUE], ws.x, nullptr, ws.x, B200_EPI_RESADD,
                          (may_defer && l + 1 < L) ? &def_f : nullptr));
        }
    }
    if (logits) {
        if (full_output) {
            RUN(launch_ln_rows((int)M, C, ws.x, 1, 0, m->gw[1], m->gw[2], ws.xo, st, pdl));
            RUN(lin(M, m->V, C, ws.xo, m->gw[3], static_cast<__half*>(logits), nullptr, nullptr, B200_EPI_NONE));
        } else {
            RUN_T(256, launch_ln_rows(B, C, ws.x, T, T - 1, m->gw[1], m->gw[2], ws.xo, st, pdl));
            RUN_T(256, lin(B, m->V, C, ws.xo, m->gw[3], static_cast<__half*>(logits), nullptr, nullptr, B200_EPI_NONE));
        }
    }
    RUN(launch_add_elapsed(B, T, elapsed, st, pdl));
#undef RUN_T
#undef RUN
    m->last_launches = n;
    return 0;
}

extern "C" int b200_layer_norm_f16(int rows, int C, const void* x, const void* w, const void* b, void* out, void* stream) {
    if (rows < 0 || C <= 0 || (C & 7) || C > 8192) return B200_ERR_BAD_ARG;
    return launch_ln_rows(rows, C, static_cast<const __half*>(x), 1, 0, static_cast<const __half*>(w),
                          static_cast<const __half*>(b), static_cast<__half*>(out), static_cast<cudaStream_t>(stream));
}

extern "C" int b200_argmax_f16(int rows, int V, const void* logits, long ld, int* out, void* stream) {
    if (rows < 0 || V <= 0 || (V & 7) || (ld & 7)) return B200_ERR_BAD_ARG;
    return launch_argmax_rows(rows, V, static_cast<const __half*>(logits), ld, out, static_cast<cudaStream_t>(stream));
}

extern "C" int b200_slot_copy(void* dst, long dst_pitch, const void* src, long src_pitch, long width_bytes, long rows, void* stream);

// One request's whole state in ONE call: the L WKV tiles, the 2L shift rows and the elapsed counter (three async copies).
// The Python-side cost of a migration (three ctypes calls + stream bookkeeping, ~100 us) was what capped round 1's
// 330-350 GB/s: the copy engine needs ~45 us for a 31.5 MiB slot.
extern "C" int b200_slot_move(void* dst1, long dst1_pitch, const void* src1, long src1_pitch, long width1, long rows1, void* dst0,
                              long dst0_pitch, const void* src0, long src0_pitch, long width0, long rows0, int* dst_elapsed,
                              const int* src_elapsed, void* stream) {
    int rc = b200_slot_copy(dst1, dst1_pitch, src1, src1_pitch, width1, rows1, stream);
    if (rc) return rc;
    rc = b200_slot_copy(dst0, dst0_pitch, src0, src0_pitch, width0, rows0, stream);
    if (rc) return rc;
    if (dst_elapsed && src_elapsed)
        rc = (int)cudaMemcpyAsync(dst_elapsed, src_elapsed, sizeof(int), cudaMemcpyDefault, static_cast<cudaStream_t>(stream));
    return rc;
}

extern "C" int b200_slot_copy(void* dst, long dst_pitch, const void* src, long src_pitch, long width_bytes, long rows,
                              void* stream) {
    if (!dst || !src || width_bytes < 0 || rows < 0 || dst_pitch < width_bytes || src_pitch < width_bytes) return B200_ERR_BAD_ARG;
    if (width_bytes == 0 || rows == 0) return 0;
    // (an SM pull kernel with 16-byte peer loads was measured against this: 268-370 GB/s alone and 76-136 GB/s under a decode
    // step vs 400 / 330-350 GB/s for the copy engine, which also costs the decode step nothing — rejected,
    // profiles/r02_migration_config5_pull_kernel_64cta_rejected.json)
    return (int)cudaMemcpy2DAsync(dst, (size_t)dst_pitch, src, (size_t)src_pitch, (size_t)width_bytes, (size_t)rows,
                                  cudaMemcpyDefault, static_cast<cudaStream_t>(stream));
}
