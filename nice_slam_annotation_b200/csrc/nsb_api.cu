// C-ABI plumbing: error reporting, decoder parameter description, pack tables, pack kernel.
#include <cstdarg>
#include <cstdio>
#include <cstring>

#include "nsb_common.cuh"

namespace nsb {

static thread_local char g_err[512] = "";

int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

static long long g_launches = 0;

int check_launch(const char* what) {
    ++g_launches;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error((int)e, "%s: %s", what, cudaGetErrorString(e));
    return 0;
}

static thread_local int g_dw_cap = 0;
void dweight_grid_cap(int ctas) { g_dw_cap = ctas; }
int dweight_grid_cap_take() { const int v = g_dw_cap; g_dw_cap = 0; return v; }
static thread_local bool g_pdl_next = false;
static int g_pdl_enabled = -1;
bool pdl_enabled() {
    if (g_pdl_enabled < 0) {
        const char* e = getenv("NSB_PDL");
        g_pdl_enabled = (e && e[0] == '0') ? 0 : 1;
    }
    return g_pdl_enabled == 1;
}
void pdl_allow_next(bool on) { g_pdl_next = on && pdl_enabled(); }
bool pdl_take() {
    const bool v = g_pdl_next;
    g_pdl_next = false;
    return v;
}

// ---- optional device-side timing of the two decoder kernels (bench.py roofline) ---------------
constexpr int TIMING_MAX = 4096;
static bool g_timing = false;
constexpr int TIMING_SLOTS = 5;      // 0 decoder forward, 1 decoder backward, 2 weight gradients, 3 exchange, 4 depth max over peers
static cudaEvent_t g_ev[TIMING_SLOTS][TIMING_MAX][2];
static int g_ev_created[TIMING_SLOTS] = {};
static int g_ev_used[TIMING_SLOTS] = {};

void timing_mark(int which, int phase, cudaStream_t s) {
    if (!g_timing) return;
    int& used = g_ev_used[which];
    if (phase == 0) {
        if (used >= TIMING_MAX) return;
        if (used >= g_ev_created[which]) {
            if (cudaEventCreate(&g_ev[which][used][0]) != cudaSuccess) return;
            if (cudaEventCreate(&g_ev[which][used][1]) != cudaSuccess) return;
            g_ev_created[which] = used + 1;
        }
        cudaEventRecord(g_ev[which][used][0], s);
    } else {
        if (used >= g_ev_created[which]) return;
        cudaEventRecord(g_ev[which][used][1], s);
        ++used;
    }
}

// ---- canonical parameters ----------------------------------------------------------------------
template <int KIND>
static int64_t param_numel(int idx) {
    using C = DecCfg<KIND>;
    if (idx < 0 || idx >= C::N_PARAM) return -1;
    if (idx < 5) return (int64_t)HID * C::in_dim(idx);           // pts_linears.i.weight [32, in]
    if (idx < 10) return HID;                                    // pts_linears.i.bias
    if (C::EMBED) {
        if (idx < 15) return (int64_t)HID * C::C_IN;             // fc_c.i.weight [32, c_dim]
        if (idx < 20) return HID;
        if (idx == 20) return (int64_t)C::NOUT * HID;            // output_linear.weight [NOUT, 32]
        if (idx == 21) return C::NOUT;
        return 3 * EMB;                                          // embedder._B [3, 93]
    }
    if (idx == 10) return HID;
    return 1;
}

static inline int32_t enc(int param, int off) { return (int32_t)((param << 20) | off); }

// Fill one packed GEMM block.  slot(j,tau) -> k index of the source matrix or -1;
// col(n,gamma) -> n index of the source matrix or -1; src(k, n) -> table entry.
template <typename SlotF, typename ColF, typename SrcF>
static void put_block(int32_t* tab, int off, int J, int NT, SlotF slot, ColF col, SrcF src) {
    for (int j = 0; j < J; ++j)
        for (int n = 0; n < NT; ++n)
            for (int lane = 0; lane < 32; ++lane) {
                const int g = lane >> 2, t = lane & 3;
                int base;
                if (NT == 1) base = off + (j * 32 + lane) * 2;
                else base = off + ((j * (NT / 2) + n / 2) * 32 + lane) * 4 + 2 * (n & 1);
                for (int q = 0; q < 2; ++q) {
                    const int k = slot(j, t + 4 * q), c = col(n, g);
                    tab[base + q] = (k < 0 || c < 0) ? -1 : src(k, c);
                }
            }
}

// tcgen05 section (nsb_layout.h TcCfg): constants and K-major operand images
template <int KIND>
static void build_tc_table(int32_t* tab) {
    using C = DecCfg<KIND>;
    using T = TcCfg<KIND>;
    if constexpr (T::ENABLED) {
        for (int a = 0; a < 3; ++a)
            for (int f = 0; f < EMB; ++f) tab[T::T_B + a * EMB_PAD + f] = enc(C::P_EMB, a * EMB + f);
        for (int i = 0; i < 5; ++i)
            for (int k = 0; k < HID; ++k) {
                tab[T::T_BIAS + i * HID + k] = enc(C::P_B + i, k);
                tab[T::T_BC + i * HID + k] = enc(C::P_BC + i, k);
            }
        for (int k = 0; k < C::NOUT; ++k) tab[T::T_BO + k] = enc(C::P_BO, k);
        for (int n = 0; n < HID; ++n)
            for (int k = 0; k < EMB; ++k) {
                tab[T::T_WE + tc_kmajor_index(n, k, 64)] = enc(C::P_W + 0, n * C::in_dim(0) + k);
                tab[T::T_WE + tc_kmajor_index(HID + n, k, 64)] = enc(C::P_W + 3, n * C::in_dim(3) + k);
            }
        for (int i = 0; i < 5; ++i)
            for (int n = 0; n < HID; ++n)
                for (int k = 0; k < T::CIN; ++k)
                    tab[T::T_WC + tc_kmajor_index(HID * i + n, k, 5 * HID)] = enc(C::P_WC + i, n * T::CIN + k);
        const int offW[5] = {-1, T::T_W1, T::T_W2, T::T_W3H, T::T_W4};
        for (int i = 1; i < 5; ++i)
            for (int n = 0; n < HID; ++n)
                for (int k = 0; k < HID; ++k)
                    tab[offW[i] + tc_kmajor_index(n, k, HID)] = enc(C::P_W + i, n * C::in_dim(i) + (i == 3 ? EMB : 0) + k);
        for (int n = 0; n < C::NOUT; ++n)
            for (int k = 0; k < HID; ++k) tab[T::T_WO + tc_kmajor_index(n, k, T::NO)] = enc(C::P_WO, n * HID + k);
        // backward blocks: rows 0..31 = W_i^T (element (n = input feature, k = output feature) = W_i[k][n]); rows
        // 32..63 are products of two parameters, written by k_pack_products after this gather
        const int offG[5] = {-1, T::T_BG1, T::T_BG2, T::T_BG3, T::T_BG4};
        for (int i = 1; i < 5; ++i)
            for (int n = 0; n < HID; ++n)
                for (int k = 0; k < HID; ++k)
                    tab[offG[i] + tc_kmajor_index(n, k, 64)] = enc(C::P_W + i, k * C::in_dim(i) + (i == 3 ? EMB : 0) + n);
        for (int n = 0; n < HID; ++n)
            for (int k = 0; k < C::NOUT; ++k) tab[T::T_BG5 + tc_kmajor_index(n, k, 64)] = enc(C::P_WO, k * HID + n);
        for (int f = 0; f < EMB; ++f)
            for (int k = 0; k < HID; ++k) {
                tab[T::T_BE0 + tc_kmajor_index(f, k, EMB_PAD)] = enc(C::P_W + 0, k * C::in_dim(0) + f);
                tab[T::T_BE3 + tc_kmajor_index(f, k, EMB_PAD)] = enc(C::P_W + 3, k * C::in_dim(3) + f);
            }
    }
}

// rows 32..63 of the backward blocks: (W_i[:, hidden] Wc_{i-1})^T for i = 1..4 and (Wo Wc_4)^T, own-grid channels
// (the first 32 columns of Wc; the fine decoder's other 32 are the middle feature, which carries no gradient).
// One block per product, one thread per element; fp32 FMA in a fixed order.
template <int KIND>
__global__ void k_pack_products(const float* const* __restrict__ ptrs, float* __restrict__ blob) {
    using C = DecCfg<KIND>;
    using T = TcCfg<KIND>;
    pdl_launch_dependents();
    pdl_wait();
    const int i = blockIdx.x + 1;                              // 1..5
    const int k = threadIdx.x >> 5, c = threadIdx.x & 31;      // k: output feature of layer i (row of W_i), c: grid channel
    const float* Wc = ptrs[C::P_WC + i - 1];                   // [32, C_IN]
    float acc = 0.f;
    if (i < 5) {
        const float* W = ptrs[C::P_W + i] + (i == 3 ? EMB : 0);
        const int ld = C::in_dim(i);
        for (int j = 0; j < HID; ++j) acc = fmaf(W[k * ld + j], Wc[j * C::C_IN + c], acc);
        const int off = i == 1 ? T::T_BG1 : (i == 2 ? T::T_BG2 : (i == 3 ? T::T_BG3 : T::T_BG4));
        blob[C::TOTAL + off + tc_kmajor_index(32 + c, k, 64)] = acc;
    } else if (k < 8) {
        if (k < C::NOUT) {
            const float* Wo = ptrs[C::P_WO];                   // [NOUT, 32]
            for (int j = 0; j < HID; ++j) acc = fmaf(Wo[k * HID + j], Wc[j * C::C_IN + c], acc);
        }
        blob[C::TOTAL + T::T_BG5 + tc_kmajor_index(32 + c, k, 64)] = acc;
    }
}

template <int KIND>
static void build_table(int32_t* tab) {
    using C = DecCfg<KIND>;
    for (int i = 0; i < BlobCfg<KIND>::TOTAL; ++i) tab[i] = -1;
    build_tc_table<KIND>(tab + C::TOTAL);
    if (C::EMBED)
        for (int a = 0; a < 3; ++a)
            for (int f = 0; f < EMB; ++f) tab[C::OFF_B + a * EMB_PAD + f] = enc(C::P_EMB, a * EMB + f);
    for (int i = 0; i < 5; ++i)
        for (int k = 0; k < HID; ++k) {
            tab[C::OFF_BIAS + i * HID + k] = enc(C::P_B + i, k);
            if (C::EMBED) tab[C::OFF_BC + i * HID + k] = enc(C::P_BC + i, k);
        }
    for (int k = 0; k < C::NOUT; ++k) tab[C::OFF_BO + k] = enc(C::P_BO, k);

    auto nat_out = [](int n, int g) { return 8 * n + g; };
    // W_i as forward operand: B[k = input feature][n = output feature] = W[n][k]
    auto fwdW = [](int i) {
        const int in = C::in_dim(i);
        return [i, in](int k, int n) { return enc(C::P_W + i, n * in + k); };
    };
    // W_i as backward operand: B[k = output feature][n = input feature] = W[k][n]
    auto bwdW = [](int i) {
        const int in = C::in_dim(i);
        return [i, in](int k, int n) { return enc(C::P_W + i, k * in + n); };
    };
    auto chain_slot = [](int j, int tau) { return slot_chain(j, tau); };
    auto gather_slot = [](int j, int tau) { return slot_gather(j, tau); };
    auto gather_col = [](int n, int g) { return col_gather(n, g); };

    const int offL[5] = {C::OFF_L0, C::OFF_L1, C::OFF_L2, C::OFF_L3, C::OFF_L4};
    const int offC[5] = {C::OFF_C0, C::OFF_C1, C::OFF_C2, C::OFF_C3, C::OFF_C4};
    const int offTL[5] = {C::OFF_TL0, C::OFF_TL1, C::OFF_TL2, C::OFF_TL3, C::OFF_TL4};
    const int offTC[5] = {C::OFF_TC0, C::OFF_TC1, C::OFF_TC2, C::OFF_TC3, C::OFF_TC4};

    if (C::EMBED) {
        auto emb_slot = [](int j, int tau) { int f = slot_natural(j, tau); return f < EMB ? f : -1; };
        auto emb_col = [](int n, int g) { int f = 8 * n + g; return f < EMB ? f : -1; };
        // forward
        put_block(tab, C::OFF_L0, E_SLABS, 4, emb_slot, nat_out, fwdW(0));
        put_block(tab, C::OFF_L1, 4, 4, chain_slot, nat_out, fwdW(1));
        put_block(tab, C::OFF_L2, 4, 4, chain_slot, nat_out, fwdW(2));
        put_block(tab, C::OFF_L3, E_SLABS, 4, emb_slot, nat_out, fwdW(3));
        put_block(tab, C::OFF_L3 + E_SLABS * C::SZ_SLAB, 4, 4,
                  [](int j, int tau) { return EMB + slot_chain(j, tau); }, nat_out, fwdW(3));
        put_block(tab, C::OFF_L4, 4, 4, chain_slot, nat_out, fwdW(4));
        for (int i = 0; i < 5; ++i)
            put_block(tab, offC[i], C::CS, 4, gather_slot, nat_out,
                      [i](int k, int n) { return enc(C::P_WC + i, n * C::C_IN + k); });
        // backward
        put_block(tab, C::OFF_TL0, 4, E_SLABS, chain_slot, emb_col, bwdW(0));
        put_block(tab, C::OFF_TL1, 4, 4, chain_slot, nat_out, bwdW(1));
        put_block(tab, C::OFF_TL2, 4, 4, chain_slot, nat_out, bwdW(2));
        put_block(tab, C::OFF_TL3X, 4, E_SLABS, chain_slot, emb_col, bwdW(3));
        put_block(tab, C::OFF_TL3, 4, 4, chain_slot, [](int n, int g) { return EMB + 8 * n + g; }, bwdW(3));
        put_block(tab, C::OFF_TL4, 4, 4, chain_slot, nat_out, bwdW(4));
        for (int i = 0; i < 5; ++i)
            put_block(tab, offTC[i], 4, 4, chain_slot, gather_col,
                      [i](int k, int n) { return enc(C::P_WC + i, k * C::C_IN + n); });
    } else {
        put_block(tab, C::OFF_L0, 4, 4, gather_slot, nat_out, fwdW(0));
        put_block(tab, C::OFF_L1, 4, 4, chain_slot, nat_out, fwdW(1));
        put_block(tab, C::OFF_L2, 4, 4, chain_slot, nat_out, fwdW(2));
        put_block(tab, C::OFF_L3, 4, 4, gather_slot, nat_out, fwdW(3));
        put_block(tab, C::OFF_L3 + 4 * C::SZ_SLAB, 4, 4,
                  [](int j, int tau) { return CDIM + slot_chain(j, tau); }, nat_out, fwdW(3));
        put_block(tab, C::OFF_L4, 4, 4, chain_slot, nat_out, fwdW(4));
        put_block(tab, C::OFF_TL0, 4, 4, chain_slot, gather_col, bwdW(0));
        put_block(tab, C::OFF_TL1, 4, 4, chain_slot, nat_out, bwdW(1));
        put_block(tab, C::OFF_TL2, 4, 4, chain_slot, nat_out, bwdW(2));
        put_block(tab, C::OFF_TL3X, 4, 4, chain_slot, gather_col, bwdW(3));
        put_block(tab, C::OFF_TL3, 4, 4, chain_slot, [](int n, int g) { return CDIM + 8 * n + g; }, bwdW(3));
        put_block(tab, C::OFF_TL4, 4, 4, chain_slot, nat_out, bwdW(4));
    }
    (void)offL; (void)offTL;
    // output layer: forward B[k = hidden][n = output] = Wo[n][k] (one n-tile);
    // backward B[k = output (natural slot)][n = hidden] = Wo[k][n]
    put_block(tab, C::OFF_OUT, 4, 1, chain_slot,
              [](int n, int g) { (void)n; return g < C::NOUT ? g : -1; },
              [](int k, int n) { return enc(C::P_WO, n * HID + k); });
    put_block(tab, C::OFF_TOUT, 1, 4,
              [](int j, int tau) { (void)j; return tau < C::NOUT ? tau : -1; }, nat_out,
              [](int k, int n) { return enc(C::P_WO, k * HID + n); });
}

__global__ void k_pack(const float* const* __restrict__ ptrs, const int32_t* __restrict__ table,
                       float* __restrict__ out, int n) {
    pdl_launch_dependents();     // the next kernel of the stream may be scheduled; it waits for this one below
    pdl_wait();                  // (programmatic launch) everything read below is complete and visible
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int32_t e = table[i];
    out[i] = e < 0 ? 0.f : ptrs[e >> 20][e & 0xfffff];
}

}  // namespace nsb

using namespace nsb;

#define KIND_SWITCH(kind, EXPR)                                            \
    switch (kind) {                                                        \
        case KIND_COARSE: { constexpr int K = KIND_COARSE; EXPR; } break;  \
        case KIND_MIDDLE: { constexpr int K = KIND_MIDDLE; EXPR; } break;  \
        case KIND_FINE: { constexpr int K = KIND_FINE; EXPR; } break;      \
        case KIND_COLOR: { constexpr int K = KIND_COLOR; EXPR; } break;    \
        default: return set_error(NSB_EINVAL, "bad decoder kind %d", kind); \
    }

extern "C" int nsb_version(void) { return NSB_VERSION; }

extern "C" int64_t nsb_launch_count(int reset) {
    long long v = g_launches;
    if (reset) g_launches = 0;
    return v;
}

extern "C" int nsb_timing(int enable) {
    g_timing = enable != 0;
    for (int w = 0; w < TIMING_SLOTS; ++w) g_ev_used[w] = 0;
    return 0;
}

extern "C" int nsb_timing_read_kernel(int which, double* ms, int* n) {
    if (which < 0 || which >= TIMING_SLOTS) return set_error(NSB_EINVAL, "bad timing slot %d", which);
    double acc = 0.0;
    for (int i = 0; i < g_ev_used[which]; ++i) {
        cudaError_t e = cudaEventSynchronize(g_ev[which][i][1]);
        float t = 0.f;
        if (e == cudaSuccess) e = cudaEventElapsedTime(&t, g_ev[which][i][0], g_ev[which][i][1]);
        if (e != cudaSuccess) return set_error((int)e, "timing: %s", cudaGetErrorString(e));
        acc += t;
    }
    if (ms) *ms = acc;
    if (n) *n = g_ev_used[which];
    return 0;
}

extern "C" int nsb_timing_read(double* ms_forward, int* n_forward, double* ms_backward, int* n_backward) {
    double ms[2] = {0.0, 0.0};
    for (int w = 0; w < 2; ++w)
        for (int i = 0; i < g_ev_used[w]; ++i) {
            cudaError_t e = cudaEventSynchronize(g_ev[w][i][1]);
            float t = 0.f;
            if (e == cudaSuccess) e = cudaEventElapsedTime(&t, g_ev[w][i][0], g_ev[w][i][1]);
            if (e != cudaSuccess) return set_error((int)e, "timing: %s", cudaGetErrorString(e));
            ms[w] += t;
        }
    if (ms_forward) *ms_forward = ms[0];
    if (n_forward) *n_forward = g_ev_used[0];
    if (ms_backward) *ms_backward = ms[1];
    if (n_backward) *n_backward = g_ev_used[1];
    return 0;
}
extern "C" const char* nsb_last_error(void) { return g_err; }

extern "C" int nsb_param_count(int kind) {
    KIND_SWITCH(kind, return DecCfg<K>::N_PARAM);
    return 0;
}

extern "C" int64_t nsb_param_numel(int kind, int index) {
    KIND_SWITCH(kind, return param_numel<K>(index));
    return 0;
}

extern "C" int64_t nsb_flat_numel(int kind) {
    int n = nsb_param_count(kind);
    if (n < 0) return n;
    int64_t s = 0;
    for (int i = 0; i < n; ++i) s += nsb_param_numel(kind, i);
    return s;
}

extern "C" int64_t nsb_packed_numel(int kind) {
    KIND_SWITCH(kind, return BlobCfg<K>::TOTAL);
    return 0;
}

extern "C" int nsb_pack_table_host(int kind, int32_t* table_host) {
    if (!table_host) return set_error(NSB_EINVAL, "null table");
    KIND_SWITCH(kind, build_table<K>(table_host));
    return 0;
}

template <int KIND>
static void layout_offsets(int32_t* o) {
    using C = DecCfg<KIND>;
    const int v[] = {C::OFF_B, C::OFF_BIAS, C::OFF_BC, C::OFF_BO,
                     C::OFF_L0, C::OFF_L1, C::OFF_L2, C::OFF_L3, C::OFF_L4,
                     C::OFF_C0, C::OFF_C1, C::OFF_C2, C::OFF_C3, C::OFF_C4, C::OFF_OUT, C::OFF_TOUT,
                     C::OFF_TL0, C::OFF_TL1, C::OFF_TL2, C::OFF_TL3, C::OFF_TL4, C::OFF_TL3X,
                     C::OFF_TC0, C::OFF_TC1, C::OFF_TC2, C::OFF_TC3, C::OFF_TC4, C::FWD_END, C::TOTAL,
                     C::CS, C::NOUT, C::EMBED ? 1 : 0};
    for (int i = 0; i < 32; ++i) o[i] = v[i];
}

extern "C" int nsb_layout_offsets(int kind, int32_t* out32_host) {
    if (!out32_host) return set_error(NSB_EINVAL, "null argument");
    KIND_SWITCH(kind, layout_offsets<K>(out32_host));
    return 0;
}

extern "C" int nsb_pack_weights(int kind, const float* const* param_ptrs, const int32_t* table,
                                float* packed, void* stream) {
    if (!param_ptrs || !table || !packed) return set_error(NSB_EINVAL, "null argument");
    int64_t n = nsb_packed_numel(kind);
    if (n < 0) return (int)n;
    launch_pdl(k_pack, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, (cudaStream_t)stream, param_ptrs, table, packed, (int)n);
    if (int rc = check_launch("k_pack")) return rc;
    switch (kind) {            // the tcgen05 backward blocks hold products of parameters (TcCfg::T_BG*)
        case KIND_MIDDLE: launch_pdl(k_pack_products<KIND_MIDDLE>, dim3(5), dim3(1024), 0, (cudaStream_t)stream, param_ptrs, packed); break;
        case KIND_FINE: launch_pdl(k_pack_products<KIND_FINE>, dim3(5), dim3(1024), 0, (cudaStream_t)stream, param_ptrs, packed); break;
        case KIND_COLOR: launch_pdl(k_pack_products<KIND_COLOR>, dim3(5), dim3(1024), 0, (cudaStream_t)stream, param_ptrs, packed); break;
        default: return 0;     // coarse: no tcgen05 section
    }
    return check_launch("k_pack_products");
}
