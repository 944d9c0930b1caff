// Packed decoder blob layout, shared by the host table generator (nsb_api.cu) and the
// device kernels (nsb_decoder.cu).  One blob per decoder:
//
//   [ B 3x96 ][ biases ][ FWD fragments ][ BWD fragments ]
//
// All GEMM operands are stored in mma.sync.m16n8k8 B-fragment order so a warp reads them
// with conflict-free 128-bit shared loads: for a GEMM with J k-slabs and NT n-tiles the
// float4 at ((j*(NT/2) + h)*32 + lane) holds (b0,b1) of n-tile 2h and (b0,b1) of n-tile
// 2h+1 for lane = 4*g + t, where b0/b1 are the k-slots t and t+4 of slab j, column g of
// the n-tile.  Which input feature a k-slot means ("slot map") and which output feature
// an accumulator column means ("column map") is chosen per GEMM so that activations never
// leave registers between layers (see nsb_decoder.cu).
#pragma once
#include <stdint.h>

#ifndef NSB_HD
#ifdef __CUDACC__
#define NSB_HD __host__ __device__
#else
#define NSB_HD
#endif
#endif

namespace nsb {

constexpr int HID = 32;        // hidden width (decoder.py:318)
constexpr int EMB = 93;        // fourier features (decoder.py:127-130)
constexpr int EMB_PAD = 96;
constexpr int E_SLABS = 12;
constexpr int CDIM = 32;       // feature channels per grid

constexpr int KIND_COARSE = 0, KIND_MIDDLE = 1, KIND_FINE = 2, KIND_COLOR = 3;

template <int KIND>
struct DecCfg {
    static constexpr bool EMBED = KIND != KIND_COARSE;
    static constexpr int CS = KIND == KIND_FINE ? 8 : 4;      // k-slabs of the feature input
    static constexpr int NOUT = KIND == KIND_COLOR ? 4 : 1;
    // ---- blob offsets in floats
    static constexpr int OFF_B = 0;
    static constexpr int SZ_B = EMBED ? 3 * EMB_PAD : 0;
    static constexpr int OFF_BIAS = OFF_B + SZ_B;
    // biases: b[5][32], (bc[5][32] if EMBED), bo[8]
    static constexpr int OFF_BC = OFF_BIAS + 5 * HID;
    static constexpr int OFF_BO = OFF_BC + (EMBED ? 5 * HID : 0);
    static constexpr int OFF_FWD = OFF_BO + 8;
    // forward GEMMs.  EMBED: L0 (12 slabs) C0 L1 C1 L2 C2 L3 (12+4 slabs) C3 L4 C4 OUT
    //                 COARSE: L0 (4, gather map) L1 L2 L3 (4 gather + 4 chain) L4 OUT
    static constexpr int SZ_SLAB = 256;                        // floats per k-slab with 4 n-tiles
    static constexpr int L0_SLABS = EMBED ? E_SLABS : 4;
    static constexpr int L3_SLABS = EMBED ? E_SLABS + 4 : 8;
    static constexpr int SZ_C = EMBED ? CS * SZ_SLAB : 0;
    static constexpr int OFF_L0 = OFF_FWD;
    static constexpr int OFF_C0 = OFF_L0 + L0_SLABS * SZ_SLAB;
    static constexpr int OFF_L1 = OFF_C0 + SZ_C;
    static constexpr int OFF_C1 = OFF_L1 + 4 * SZ_SLAB;
    static constexpr int OFF_L2 = OFF_C1 + SZ_C;
    static constexpr int OFF_C2 = OFF_L2 + 4 * SZ_SLAB;
    static constexpr int OFF_L3 = OFF_C2 + SZ_C;
    static constexpr int OFF_C3 = OFF_L3 + L3_SLABS * SZ_SLAB;
    static constexpr int OFF_L4 = OFF_C3 + SZ_C;
    static constexpr int OFF_C4 = OFF_L4 + 4 * SZ_SLAB;
    static constexpr int OFF_OUT = OFF_C4 + SZ_C;              // 4 slabs x 32 lanes x float2
    static constexpr int OFF_BWD = OFF_OUT + 4 * 64;
    static constexpr int FWD_END = OFF_BWD;
    // backward (transposed) GEMMs, K = 32 (4 slabs) except TOUT (1 slab, K = 8 padded).
    //   TOUT : N = 32                       d(out) -> d(h5)
    //   TLi  : N = 32                       d(a_i) -> d(h_i)      natural column map
    //   TCi  : N = 32                       d(h_{i+1}) -> d(c)    gather column map
    //   TL3X / TL0 : the input-side part of layers 3 / 0:  EMBED d(a) -> d(e), N = 96,
    //   natural map; COARSE d(a) -> d(c), N = 32, gather map.  TL3 is the h part of layer 3.
    static constexpr int X_NT = EMBED ? 12 : 4;                // n-tiles of the TL3X / TL0 blocks
    static constexpr int SZ_TC = EMBED ? 4 * SZ_SLAB : 0;
    static constexpr int OFF_TOUT = OFF_BWD;                   // 1 slab x 4 n-tiles
    static constexpr int OFF_TL4 = OFF_TOUT + SZ_SLAB;
    static constexpr int OFF_TC4 = OFF_TL4 + 4 * SZ_SLAB;
    static constexpr int OFF_TL3 = OFF_TC4 + SZ_TC;
    static constexpr int OFF_TL3X = OFF_TL3 + 4 * SZ_SLAB;
    static constexpr int OFF_TC3 = OFF_TL3X + 4 * X_NT * 64;
    static constexpr int OFF_TL2 = OFF_TC3 + SZ_TC;
    static constexpr int OFF_TC2 = OFF_TL2 + 4 * SZ_SLAB;
    static constexpr int OFF_TL1 = OFF_TC2 + SZ_TC;
    static constexpr int OFF_TC1 = OFF_TL1 + 4 * SZ_SLAB;
    static constexpr int OFF_TL0 = OFF_TC1 + SZ_TC;
    static constexpr int OFF_TC0 = OFF_TL0 + 4 * X_NT * 64;
    static constexpr int TOTAL = OFF_TC0 + SZ_TC;
    // ---- canonical parameters
    static constexpr int N_PARAM = EMBED ? 23 : 12;
    static constexpr int P_W = 0;                              // pts_linears.i.weight
    static constexpr int P_B = 5;                              // pts_linears.i.bias
    static constexpr int P_WC = 10;                            // fc_c.i.weight (EMBED)
    static constexpr int P_BC = 15;
    static constexpr int P_WO = EMBED ? 20 : 10;
    static constexpr int P_BO = EMBED ? 21 : 11;
    static constexpr int P_EMB = 22;
    static constexpr int C_IN = KIND == KIND_FINE ? 64 : 32;   // fc_c input width
    NSB_HD static constexpr int in_dim(int i) {
        return EMBED ? (i == 0 ? EMB : (i == 3 ? EMB + HID : HID)) : (i == 3 ? 2 * HID : HID);
    }
};

// slot maps: feature index of k-slot tau (0..7) of slab j
NSB_HD inline int slot_natural(int j, int tau) { return 8 * j + tau; }
NSB_HD inline int slot_chain(int j, int tau) { return 8 * j + 2 * (tau & 3) + (tau >> 2); }
// gather map: lane t holds channels 8t..8t+7 of its rows; slab j (mod 4) uses 8t+2j, 8t+2j+1;
// slabs 4..7 (fine decoder) are the concatenated middle-grid feature (channels 32..63)
NSB_HD inline int slot_gather(int j, int tau) {
    return 32 * (j >> 2) + 8 * (tau & 3) + 2 * (j & 3) + (tau >> 2);
}
// column maps: output feature of accumulator column gamma (0..7) of n-tile n
NSB_HD inline int col_natural(int n, int gamma) { return 8 * n + gamma; }
NSB_HD inline int col_gather(int n, int gamma) { return 8 * (gamma >> 1) + 2 * n + (gamma & 1); }

}  // namespace nsb

// ---- tcgen05 blob ------------------------------------------------------------------------------
// Second section of the packed blob (after DecCfg::TOTAL floats), read by the tcgen05 kernels
// (nsb_decoder_tc.cu).  Weight blocks are stored as the UMMA "K-major, no swizzle" operand
// image of an R-row matrix: element (r, k) at float index
//   ((k / 4) * (R / 8) + r / 8) * 32 + (r % 8) * 4 + (k % 4)          (nsb_tc.cuh kmajor_index)
// so that a CTA copies them to shared memory verbatim (splitting every value into its tf32
// hi / lo parts on the way) and addresses them with shared-memory matrix descriptors.
namespace nsb {
NSB_HD inline int tc_kmajor_index(int r, int k, int R) {
    return ((k >> 2) * (R >> 3) + (r >> 3)) * 32 + (r & 7) * 4 + (k & 3);
}
template <int KIND>
struct TcCfg {
    using C = DecCfg<KIND>;
    static constexpr bool ENABLED = C::EMBED;                  // coarse (MLP_no_xyz) stays on mma.sync
    static constexpr int CIN = C::C_IN;
    static constexpr int NO = 16;                              // padded output rows (UMMA N >= 16 at M = 128)
    // constants: B [3][96], b [5][32], bc [5][32], bo [16]
    static constexpr int T_B = 0;
    static constexpr int T_BIAS = T_B + 3 * EMB_PAD;
    static constexpr int T_BC = T_BIAS + 5 * HID;
    static constexpr int T_BO = T_BC + 5 * HID;
    static constexpr int T_CONST_END = T_BO + NO;              // 624
    // forward weights (operand B of D = A W^T): R = output features, K = input features
    static constexpr int T_WE = T_CONST_END;                   // R 64 : [W0 ; W3[:, :93]],  K 96
    static constexpr int T_WC = T_WE + 64 * EMB_PAD;           // R 160: [Wc0 ; .. ; Wc4],   K CIN
    static constexpr int T_W1 = T_WC + 5 * HID * CIN;          // R 32, K 32
    static constexpr int T_W2 = T_W1 + HID * HID;
    static constexpr int T_W3H = T_W2 + HID * HID;             // W3[:, 93:]
    static constexpr int T_W4 = T_W3H + HID * HID;
    static constexpr int T_WO = T_W4 + HID * HID;              // R 16 (rows >= NOUT zero), K 32
    static constexpr int T_FWD_END = T_WO + NO * HID;
    static constexpr int W_FWD = T_FWD_END - T_CONST_END;      // forward weight floats
    // backward operands (k_decoders_backward_tc, nsb_decoder_bwd_tc.cuh).  With da_i = d(a_i) (gradient of the
    // pre-activation of layer i, da_5 := d(out)) the adjoint chain is  dh_{i-1}' = da_i W_i[:, hidden part]  and the
    // feature gradient  dc = sum_i dh_i' Wc_{i-1} = sum_i da_i (W_i[:, hidden] Wc_{i-1}):  ONE operand block per
    // layer, R = 64 rows = [W_i^T (32 input features) ; (W_i Wc_{i-1})^T (32 own-grid channels)], K = 32 (layer
    // outputs; block 5: K = 8 padded outputs of the head), so a layer is one N = 64 MMA group with A = da_i.  The
    // product rows are filled by k_pack_products after the gather.  T_BE0 / T_BE3: W_0[:, :93]^T, W_3[:, :93]^T
    // (R = 96) for d(e) -> d(point), d(B).
    static constexpr int T_BG1 = T_FWD_END;                    // R 64, K 32: [W_1^T ; (W_1 Wc_0)^T]
    static constexpr int T_BG2 = T_BG1 + 64 * HID;
    static constexpr int T_BG3 = T_BG2 + 64 * HID;             // W_3[:, 93:]
    static constexpr int T_BG4 = T_BG3 + 64 * HID;
    static constexpr int T_BG5 = T_BG4 + 64 * HID;             // R 64, K 8: [Wo^T ; (Wo Wc_4)^T], outputs >= NOUT zero
    static constexpr int T_BE0 = T_BG5 + 64 * 8;               // R 96, K 32
    static constexpr int T_BE3 = T_BE0 + EMB_PAD * HID;
    static constexpr int T_BWD_END = T_BE3 + EMB_PAD * HID;
    static constexpr int W_BWD = T_BWD_END - T_FWD_END;        // backward weight floats (14848)
    static constexpr int TOTAL = ENABLED ? T_BWD_END : 0;
};
template <int KIND>
struct BlobCfg { static constexpr int TOTAL = DecCfg<KIND>::TOTAL + TcCfg<KIND>::TOTAL; };
}  // namespace nsb

// ---- weight-gradient scratch record ---------------------------------------------------------
// One record per 16-point tile, written by the forward (layer inputs X) and the backward kernel (output-side
// gradients G) and consumed by k_dweight / k_dweight_tc: d(W) of a layer is G^T X summed over points.
// Every 32-channel block of a section is an "atom": a dense [16 points][32 channels] fp32 matrix (128 bytes per
// point) whose 32-byte chunks are XOR-swizzled with (point & 3) (rec_sw).  That is at once bank-conflict free for
// k_dweight's transposed mma.sync fragment reads and the ONLY shared-memory layout tcgen05 accepts for MN-major
// tf32 operands (SWIZZLE_128B_BASE32B: 4-point x 128-byte swizzle atoms, next 4 points 512 B further = SBO, next
// 32 channels 2048 B further = LBO; scripts/probe_mn_decode.cu), so a record copied verbatim into shared memory is
// the operand image of k_dweight_tc.  Atoms that one MMA stacks along M or N are adjacent:
//   [c | e0 e1 e2]           X stack of W0 / W3 (embedding part) / Wc_i
//   [h1 h2 h3 h4] h5         X stack of W1..W4; h5 feeds the output layer
//   [da1 da2 da4 da3] da0    G of [h1 h2 h4 h3]; the last two ...
//   ... [da3 da0 dh0..dh4]   are the first two of the G stack of [c | e]
namespace nsb {
NSB_HD inline int rec_da_slot(int i) { return i == 0 ? 4 : (i == 3 ? 3 : (i == 4 ? 2 : i - 1)); }
template <int KIND>
struct RecCfg {
    using C = DecCfg<KIND>;
    static constexpr int ATOM = 16 * HID;                      // floats of one atom (2048 B)
    static constexpr int W_E = C::EMBED ? EMB_PAD : 0;
    static constexpr int R_C = 0;                              // c                    [16][C_IN]
    static constexpr int R_E = R_C + 16 * C::C_IN;             // e = sin(pB)          3 atoms (features 32j..32j+31)
    static constexpr int R_H = R_E + 16 * W_E;                 // h1..h5               5 atoms
    static constexpr int R_P = R_H + 5 * ATOM;                 // p (cols 0..2)        [16][8]
    static constexpr int R_DA = R_P + 16 * 8;                  // d(a_i) at slot rec_da_slot(i)   5 atoms
    static constexpr int R_DH = R_DA + 5 * ATOM;               // d(h1)..d(h5)         5 atoms
    static constexpr int R_DO = R_DH + 5 * ATOM;               // d(out) (cols < NOUT) 1 atom
    static constexpr int R_DE = R_DO + ATOM;                   // d(e) * cos(pB)       3 atoms
    static constexpr int REC = R_DE + 16 * W_E;
    NSB_HD static int da(int i) { return R_DA + rec_da_slot(i) * ATOM; }
};
constexpr int REC_MAX = RecCfg<KIND_FINE>::REC;
// column of element (row, col) inside a row of a section of width >= 32 (see above)
NSB_HD inline int rec_sw(int row, int col) { return col ^ ((row & 3) << 3); }
// float offset of feature f of point `row` inside an e / d(e) section (3 atoms)
NSB_HD inline int rec_e(int row, int f) { return (f >> 5) * (16 * HID) + row * HID + rec_sw(row, f & 31); }
}  // namespace nsb
