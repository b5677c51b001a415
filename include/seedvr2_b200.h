/*
 * seedvr2_b200.h — C ABI of libseedvr2_b200.so: the B200-native (sm_100a) hot path of SeedVR2's one-step NaDiT
 * transformer block.  Every entry point takes raw device pointers, sizes and a cudaStream_t (passed as void*), is an
 * asynchronous stream-ordered launch, never allocates or retains caller memory and returns 0 on success or a negative
 * SVR2_E_* code (message via svr2_last_error()).  No torch types appear here; PyTorch only owns the memory.
 *
 * Each function cites the reference code (disk02/ComfyUI-SeedVR2 v2.5.18, paths relative to its root) it replaces.
 * The reference is pure Python, so "the reference FFI for this path" is a ctypes binding: see INTEGRATION.md.
 */
#ifndef SEEDVR2_B200_H
#define SEEDVR2_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SVR2_OK             0
#define SVR2_E_INVALID     -1   /* bad argument (null pointer, unsupported size, b != 1, head_dim != 128 ...) */
#define SVR2_E_CUDA        -2   /* CUDA runtime / driver error, text in svr2_last_error() */
#define SVR2_E_UNSUPPORTED -3   /* valid request this build does not implement (never a silent fallback) */

#define SVR2_ROPE_NONE   0      /* 7B text rows */
#define SVR2_ROPE_3B_MM  1      /* dit_3b/rope.py:88-176  lang freqs, positions (Lt+t, h, w) / text (j,j,j) */
#define SVR2_ROPE_7B_PIX 2      /* dit_7b/rope.py:73-111  pixel freqs, linspace(-1,1,n) per window, bf16 angles, video only */

#define SVR2_HEAD_DIM 128
#define SVR2_WIN_FIELDS 8       /* int32 per window in a window table: t0,h0,w0, nt,nh,nw, row_start, n_vid */

/* Last error text of the calling thread ("" if none). */
const char* svr2_last_error(void);
/* ABI version (bumped when a signature changes). */
int svr2_abi_version(void);

/*
 * Window geometry, host side, pure integer/double arithmetic.
 * Replaces make_720Pwindows_bysize / make_shifted_720Pwindows_bysize (src/models/dit_3b/window.py:28-83, same file in
 * dit_7b) incl. Python banker's round, ceil of double quotients and int() truncation.
 *   T,H,W        token grid;  nt,nh,nw  the YAML `window` triple (4,3,3);  shifted  0/1;  txt_len  Lt
 *   win_table    [max_windows][SVR2_WIN_FIELDS] int32, windows in the reference order (iw outer, ih, it inner);
 *                row_start = first packed row of the window's sequence [vid rows ; txt rows]
 *   out_counts   [0]=n_windows [1]=sum Lw (== T*H*W) [2]=max Lw [3]=packed rows P = sum Lw + n_windows*Lt
 *                [4]=number of 128-row query tiles over all windows [5]=min Lw
 * Returns SVR2_E_INVALID if max_windows is too small (out_counts[0] is still set).
 */
int svr2_window_plan(int T, int H, int W, int nt, int nh, int nw, int shifted, int txt_len,
                     int32_t* win_table, int max_windows, int64_t* out_counts);

/*
 * Window partition indices on the device (bit-exact integers).
 * Replaces na.window_idx / na.window (src/models/dit_3b/na.py:583-641): tgt_idx (gather), src_idx (its inverse, the
 * reference's argsort), window_shape [nw,3] int64, cu_seqlens [nw+1] int32 with Lt text rows appended per window
 * (mmattn.py:211-255).  win_table is the DEVICE copy of svr2_window_plan's table.
 */
int svr2_window_index(const int32_t* win_table, int n_windows, int T, int H, int W, int txt_len,
                      int64_t* tgt_idx, int64_t* src_idx, int64_t* window_shape, int32_t* cu_seqlens, void* stream);

/*
 * Fused RMSNorm (no affine) + adaLN "in" modulation:  y = rmsnorm(x) * (scaleA + scaleB) + (shiftA + shiftB)
 * Replaces CustomRMSNorm.forward (src/models/dit_3b/normalization.py:88-109) + AdaSingle.forward mode="in"
 * (modulation.py:65-110) — about nine ATen kernels.  bf16 in/out, every intermediate rounded to bf16 where the
 * reference's op-by-op bf16 path rounds.
 *   x,y [L,D] bf16 (may alias); emb [6*D] bf16 laid out (d, layer, {shift,scale,gate}); shiftB/scaleB [D] bf16 or both
 *   NULL (norm only: the 3B last layer's vid_only text branch); layer 0=attn 1=mlp.  D % 8 == 0, D <= 4096.
 */
int svr2_modnorm_fwd(const void* x, const void* emb, const void* shiftB, const void* scaleB, void* y,
                     int64_t L, int D, int layer, float eps, void* stream);

/*
 * RMSNorm WITH elementwise-affine weight + adaLN "in" modulation:  y = rmsnorm(x) * weight * (scaleA + scaleB) + (shiftA + shiftB)
 * Replaces the model's output norm + output modulation: `vid_out_norm` (CustomRMSNorm with weight, normalization.py:88-109)
 * followed by `vid_out_ada(..., layer="out", mode="in")` (src/models/dit_3b/nadit.py:233-244).  Cache-key quirk Q1: the
 * reference's cache hands vid_out_ada the ATTN-layer (shift, scale) of emb, so the caller passes layer 0.
 *   weight [D] bf16; the rest as svr2_modnorm_fwd.
 */
int svr2_modnorm_affine_fwd(const void* x, const void* weight, const void* emb, const void* shiftB, const void* scaleB,
                            void* y, int64_t L, int D, int layer, float eps, void* stream);

/*
 * Gate + residual:  out = h * (gateA + gateB) + resid      (modulation.py:111-116 + mmsr_block.py:114,126)
 * gateB NULL with emb NULL means plain residual add (vid_only text branch).  h/resid/out [L,D] bf16, out may alias h.
 */
int svr2_gate_residual_fwd(const void* h, const void* resid, const void* emb, const void* gateB, void* out,
                           int64_t L, int D, int layer, void* stream);

/*
 * Gate + residual fused with the FOLLOWING RMSNorm + modulation (saves one [L,D] read):
 *   out = h * (gateA[layer] + gateB) + resid ;  y = rmsnorm(out) * (scaleA[layer+1] + scaleB2) + (shiftA[layer+1] + shiftB2)
 * i.e. mmsr_block.py:113-121 in one pass.  `layer` is the gate's layer (0); the modulation uses layer+1.
 */
int svr2_gate_residual_modnorm_fwd(const void* h, const void* resid, const void* emb, const void* gateB,
                                   const void* shiftB2, const void* scaleB2, void* out, void* y,
                                   int64_t L, int D, int layer, float eps, void* stream);

/*
 * SwiGLU activation:  out = silu(gate) * up, silu rounded to bf16 before the product.
 * Replaces `F.silu(self.proj_in_gate(x)) * self.proj_in(x)` (src/models/dit_3b/mlp.py:60-62) — two ATen kernels.
 * gate/up/out bf16, n elements (multiple of 8), out may alias gate or up.
 */
int svr2_swiglu_fwd(const void* gate, const void* up, void* out, int64_t n, void* stream);
/* The same on a [rows, 2*hidden] buffer whose rows hold gate | up side by side: the output of ONE GEMM with the two
 * in-projection weights of SwiGLUMLP (mlp.py:56-58) stacked.  out [rows, hidden]. */
int svr2_swiglu_rows_fwd(const void* gate_up, void* out, int64_t rows, int hidden, void* stream);

/*
 * The block's linear layers: ONE persistent tcgen05 GEMM kernel (CTA pairs, TMA-staged operands, TMEM
 * accumulators, csrc/gemm.cu) whose epilogue applies the row-wise work that follows the linear in the reference.
 * x [M, K] bf16 row-major; weights in nn.Linear layout [N, K] bf16 row-major; bias [N] bf16 or NULL; y / out [M, N] bf16.
 * N % 64 == 0, K % 8 == 0.  The accumulator (+ bias) is rounded to bf16 first — the linear's output tensor in the
 * reference — and every following op rounds again like the reference's bf16 path.
 *
 * svr2_linear_fwd: y = x W^T + bias, optionally followed by GELU-tanh.
 *   Replaces proj_qkv (src/models/dit_3b/nablocks/attention/mmattn.py:56-59,173; no bias) and, with
 *   SVR2_ACT_GELU_TANH, the 7B MLP's proj_in + F.gelu(approximate="tanh") (src/models/dit_3b/mlp.py:28-43).
 */
#define SVR2_ACT_NONE      0
#define SVR2_ACT_GELU_TANH 1
int svr2_linear_fwd(const void* x, const void* weight, const void* bias, void* y, int64_t M, int N, int K, int activation,
                    void* stream);
/*
 * svr2_linear_swiglu_fwd: y = silu(x Wg^T) * (x Wu^T) — both in-projections of SwiGLUMLP in one tile (two B operands), the
 * activation in the epilogue; the two [M, hidden] projections never reach HBM.
 *   Replaces `F.silu(self.proj_in_gate(x)) * self.proj_in(x)` (src/models/dit_3b/mlp.py:46-62).  N = hidden, N % 128 == 0.
 */
int svr2_linear_swiglu_fwd(const void* x, const void* w_gate, const void* w_up, void* y, int64_t M, int N, int K, void* stream);
/*
 * svr2_linear_gate_residual_fwd: out = (x W^T + bias) * (gateA[layer] + gateB) + resid.
 *   Replaces attn.proj_out / mlp.proj_out (mmattn.py:269, mlp.py:43,62) + AdaSingle mode="out" (modulation.py:111-116) +
 *   the residual add (mmsr_block.py:114,126).  emb [6 N] bf16 (d, layer, {shift, scale, gate}); gateB [N]; resid [M, N];
 *   out must not alias x; N <= 4096.
 */
int svr2_linear_gate_residual_fwd(const void* x, const void* weight, const void* bias, const void* emb, const void* gateB,
                                  int layer, const void* resid, void* out, int64_t M, int N, int K, void* stream);

/*
 * q/k RMSNorm (affine, per head) + 3D RoPE + window partition + text interleave, one pass over qkv.
 * Replaces window_partition index_select (na.py:637, mmattn.py:199), norm_q/norm_k (mmattn.py:207-208,
 * normalization.py:88-109), the text repeat (mmattn.py:223-236), RoPE (dit_3b/rope.py:92-127, dit_7b/rope.py:73-91 +
 * third-party apply_rotary_emb) and concat_win (na.py:320-424) — writes q,k,v directly in the packed varlen layout
 * [P, heads, 128]: sequence k = [video rows of window k ; all Lt text rows].
 *   qkv_vid [L, 3, heads, 128] bf16 (token order), qkv_txt [Lt, 3, heads, 128] bf16
 *   wq_vid,wk_vid,wq_txt,wk_txt [128] bf16;  freqs fp32 [n_freqs] = the checkpoint buffer attn.rope.rope.freqs
 *   row_map: svr2_pack_rowmap's table for this plan, or NULL (every CTA then finds its rows' windows itself — slower)
 */
int svr2_qkv_pack_fwd(const void* qkv_vid, const void* qkv_txt, const void* wq_vid, const void* wk_vid,
                      const void* wq_txt, const void* wk_txt, const float* freqs, int n_freqs, int rope_kind,
                      const int32_t* win_table, const int32_t* row_map, int n_windows, int T, int H, int W, int txt_len,
                      int heads, float eps, void* q_pack, void* k_pack, void* v_pack, int64_t P, void* stream);

/*
 * Packed-row map of a window plan, computed once per plan on the device (pure integer): for packed row r
 *   row_map[r] = { source (video token index >= 0, or -(1 + text row)), local t | local h << 16, local w, window }.
 * It is the per-row form of na.window_idx (na.py:616-641) + the text repeat order (mmattn.py:223-236); svr2_qkv_pack_fwd
 * reads it instead of searching the window table in every CTA.   row_map [P][4] int32 on the device.
 */
int svr2_pack_rowmap(const int32_t* win_table, int n_windows, int T, int H, int W, int txt_len, int32_t* row_map,
                     int64_t P, void* stream);

/*
 * The block's qkv projection with everything that follows it up to the attention kernel fused into the GEMM epilogue:
 *   q,k,v_pack = window_partition(rope(rmsnorm(x Wqkv^T)))
 * Replaces proj_qkv (mmattn.py:56-59,173 / dit_7b mmsr_block.py:85-90) + everything svr2_qkv_pack_fwd replaces (norm_q / norm_k,
 * RoPE, window_partition, concat_win; mmattn.py:199-248) — the [rows, 3*heads*128] qkv tensor never reaches HBM (SURVEY §8f N1c,
 * §8d: 0 algorithmic bytes for partition / interleave).  tcgen05 GEMM of csrc/gemm.cu; a 256-column tile is two whole heads.
 *   x [M, K] bf16 rows in this order: video tokens (original order), then text tokens — any contiguous range of that order:
 *   tok_map points at the entry of row 0 (video GEMM: tok_map, M = L; text GEMM: tok_map + 4*L, M = Lt; joined rows: M = L + Lt);
 *   weight [3*heads*128, K] bf16 (nn.Linear layout, rows ordered q heads | k heads | v heads);
 *   tok_map / rope_cs / rope_rows from svr2_rope_plan;  n_freqs 21 (3B), 10 (7B) or 0 (no rope);
 *   wq_*, wk_* [128] bf16 norm weights (the _txt pair may be NULL when no row is a text row);
 *   q,k,v_pack [P, heads, 128] bf16: video rows land at their packed position;
 *   txt_qkv [Lt, 3, heads, 128] bf16: text rows, copied into every window's sequence by svr2_txt_replicate_fwd afterwards.
 */
int svr2_linear_qkv_pack_fwd(const void* x, const void* weight, int64_t M, int K, int heads, const int32_t* tok_map,
                             const float* rope_cs, int rope_rows, int n_freqs, const void* wq_vid, const void* wk_vid,
                             const void* wq_txt, const void* wk_txt, float eps, void* q_pack, void* k_pack, void* v_pack,
                             void* txt_qkv, void* stream);

/*
 * Per-plan tables of svr2_linear_qkv_pack_fwd, computed once per (plan, rope kind, frequencies) on the device:
 *   tok_map [L + Lt][4] int32: the inverse of the packed-row map — for video token i (original order) its packed row and the
 *   rope-table row of each axis; for text row j (entry L + j): -(1 + j) and, 3B, row j for all axes (rope.py:173), else -1.
 *   rope_cs [rope_rows][n_freqs][2] fp32 (cos, sin): one row per distinct position value.  3B (dit_3b/rope.py:154-176): row r =
 *   position r, rope_rows > max(Lt + window nt, nh, nw, Lt) - 1.  7B (dit_7b/rope.py:73-111): row base_of_n[n] + i =
 *   linspace(-1, 1, n)[i] for every window extent n of the plan; base_of_n [65] int32 on the device, -1 where n is absent.
 * Same cos / sin bits as svr2_qkv_pack_fwd computes per row.  row_map from svr2_pack_rowmap.
 */
int svr2_rope_plan(const int32_t* win_table, const int32_t* row_map, int n_windows, int T, int H, int W, int txt_len, int64_t P,
                   int rope_kind, const float* freqs, int n_freqs, const int32_t* base_of_n, int rope_rows, int32_t* tok_map,
                   float* rope_cs, void* stream);

/*
 * Text rows of q, k, v into every window's sequence: q_pack[row_start(w) + n_vid(w) + j] = txt_qkv[j, 0] (k: [j, 1], v: [j, 2]).
 * Replaces the per-window repeat of the text q/k/v (mmattn.py:223-236) for the fused projection above.
 */
int svr2_txt_replicate_fwd(const void* txt_qkv, const int32_t* win_table, int n_windows, int txt_len, int heads, void* q_pack,
                           void* k_pack, void* v_pack, void* stream);

/*
 * Work list for svr2_window_attn_fwd, host side, pure integer: one item = (window, pair of 128-row query tiles), items
 * sorted longest sequence first so the ragged shifted-window remnants fill the tail of the last wave.
 *   work [max_items][2] int32 (window, q_pair) or NULL to only count;  *n_items receives the item count.
 */
int svr2_window_attn_worklist(const int32_t* win_table_host, int n_windows, int txt_len, int32_t* work,
                              int max_items, int64_t* n_items);
/* The same list in units of `tiles_per_item` query tiles (2 = the list above; 4 = the list of the CTA-pair kernel,
 * window_attn_v7.cu, whose cluster takes four query tiles of a window at a time):  work [max_items][2] = (window, group). */
int svr2_window_attn_worklist_tiles(const int32_t* win_table_host, int n_windows, int txt_len, int tiles_per_item,
                                    int32_t* work, int max_items, int64_t* n_items);

/*
 * Variable-length window attention over the packed layout, window reverse and text pooling fused in the epilogue.
 * Replaces FlashAttentionVarlen / pytorch_varlen_attention's per-window SDPA loop (src/models/dit_3b/attention.py:
 * 24-61,138-157), unconcat_win (na.py:396-417) and window_reverse (na.py:638, mmattn.py:264).
 * softmax(QK^T / sqrt(128)) V per sequence, non-causal, fp32 softmax, bf16 P.  TMA + tcgen05 + TMEM (sm_100a).
 *   q,k,v_pack [P, heads, 128] bf16;  win_table / work / work4: DEVICE copies of the plan's tables (work4 = the
 *   4-tile work list, or NULL with n_work4 = 0: the kernels that need it are then not dispatched);
 *   out_vid [L, heads*128] bf16 in ORIGINAL token order;
 *   txt_part [n_windows, Lt, heads*128] bf16 workspace (per-window text rows);  out_txt [Lt, heads*128] bf16 = mean
 *   over windows (na.py:414), or NULL to leave the pooling to svr2_txt_pool_fwd.
 */
int svr2_window_attn_fwd(const void* q_pack, const void* k_pack, const void* v_pack, const int32_t* win_table,
                         const int32_t* work, int n_work, const int32_t* work4, int n_work4, int n_windows, int T, int H,
                         int W, int txt_len, int heads, int64_t P, void* out_vid, void* txt_part, void* out_txt,
                         void* stream);

/*
 * The same call with the kernel named explicitly (tests and A/B runs; the product path calls svr2_window_attn_fwd):
 *   SVR2_ATTN_AUTO  what svr2_window_attn_fwd picks from the packed lengths
 *   SVR2_ATTN_GEN8  CTA pairs, tcgen05.mma.cta_group::2 (window_attn_v6.cu);  SVR2_ATTN_GEN4  persistent CTA per SM
 *   (window_attn_v2.cu) — the two product kernels, bit-identical outputs
 *   SVR2_ATTN_MMA   legacy mma.sync kernel (window_attn_mma.cu): an independent CUDA implementation kept as a test
 *   comparator, never selected by SVR2_ATTN_AUTO
 * Other codes name the earlier generations under csrc/experiments/ and return SVR2_E_UNSUPPORTED unless the library was
 * built with `make EXPERIMENTS=1`.
 */
#define SVR2_ATTN_AUTO 0
#define SVR2_ATTN_GEN4 4
#define SVR2_ATTN_GEN8 8
#define SVR2_ATTN_GEN10 10   /* generation 4's kernel with two threads per query row (window_attn_v2.cu, HALVES = 2) */
#define SVR2_ATTN_MMA  100
int svr2_window_attn_fwd_impl(const void* q_pack, const void* k_pack, const void* v_pack, const int32_t* win_table,
                              const int32_t* work, int n_work, const int32_t* work4, int n_work4, int n_windows, int T,
                              int H, int W, int txt_len, int heads, int64_t P, void* out_vid, void* txt_part,
                              void* out_txt, int impl, void* stream);

/*
 * svr2_window_attn_fwd without caller scratch, for callers that cannot allocate (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same inputs, outputs and kernels as svr2_window_attn_fwd; the per-window text rows (`txt_part`) live in a file-scope
 * static device array of SVR2_NOALLOC_SCRATCH_BYTES (32 MiB) that every GPU loading the library holds until exit, and the
 * text pooling always runs inside the call (out_txt is required when txt_len > 0).  The call allocates nothing (no
 * cudaMalloc / cudaFree / allocating library routine).  Limit: n_windows * txt_len * heads * 128 * 2 bytes must fit the
 * array (58 text rows: up to 112 windows at 20 heads, 94 at 24 heads) — larger problems return SVR2_E_INVALID with a message.
 * All calls on one device share the array: calls that overlap on the GPU race; stream-ordered calls are safe (nothing in
 * the array needs to start at zero, every slot that is pooled is written first).
 */
#define SVR2_NOALLOC_SCRATCH_BYTES (32u << 20)
int svr2_window_attn_fwd_noalloc(const void* q_pack, const void* k_pack, const void* v_pack, const int32_t* win_table,
                                 const int32_t* work, int n_work, int n_windows, int T, int H, int W, int txt_len, int heads,
                                 int64_t P, void* out_vid, void* out_txt, void* stream);

/*
 * Debug hook (no reference counterpart): point the attention kernel at a device buffer of n_ctas * 64 int64 slots and
 * every CTA records clock64 stamps of its pipeline phases (slot map in csrc/window_attn.cu).  NULL switches it off.
 * Returns the number of slots per CTA.  Used by scripts/attn_trace.py to read the per-CTA timeline.
 */
int svr2_debug_attn_trace(void* trace_buf, int n_ctas);

/*
 * Text pooling on its own: out_txt[j] = mean over windows of txt_part[w, j]  (na.py:414 — unconcat_coalesce's
 * `.mean(1)` over the per-window copies of the text rows; bf16 rows, fp32 accumulate, bf16 result).
 * svr2_window_attn_fwd runs it itself unless out_txt is NULL.
 */
int svr2_txt_pool_fwd(const void* txt_part, void* out_txt, int n_windows, int txt_len, int heads, void* stream);

/*
 * Overlap crossfade between neighbouring temporal segments (Hann for overlap >= 3, linear below), fp32.
 * Replaces blend_overlapping_frames (src/core/generation_utils.py:284-312) as used by the multi-GPU parent
 * (inference_cli.py:1163-1198):  out[f] = prev_tail[f]*w[f] + cur_head[f]*(1-w[f]).   out may alias either input.
 */
int svr2_halo_blend(const float* prev_tail, const float* cur_head, float* out, int overlap, int64_t frame_elems,
                    void* stream);

/*
 * Halo exchange + crossfade + gather of one rank's temporal segment in ONE kernel over peer memory (NVLink loads / stores).
 * Replaces, for this rank, the parent process's stitch loop (inference_cli.py:1163-1198) and the shared-memory / queue
 * hand-over of the worker results (inference_cli.py:1021,1148-1153).
 *   mine       this rank's segment [frames, frame_elems] fp32 (local)
 *   prev_tail  the previous rank's LAST `head` frames (peer mapping of its segment buffer), NULL when head == 0
 *   out        this rank's first output frame inside the root's stitched video (peer mapping unless this rank is the root)
 *   out[f] = prev_tail[f]*w[f] + mine[f]*(1-w[f])  for f < head  (w as svr2_halo_blend with overlap = head),
 *   out[head + g] = mine[first_body + g]            for g < n_body.
 * The caller orders the launch after every rank finished producing its segment and tells the root when all ranks are done
 * (a barrier on either side; comfyui-seedvr2_b200/segments.py uses the symmetric-memory signal pads).
 */
int svr2_segment_stitch_peer(const float* mine, const float* prev_tail, float* out, int head, int first_body, int n_body,
                             int64_t frame_elems, void* stream);

/*
 * DiT input of the one-step super-resolution call, bf16:  out[n, 0:C] = noise[n], out[n, C:2C] = blurred latent[n],
 * out[n, 2C] = 1.   Replaces get_condition(task="sr") (src/core/infer.py:54-78) + torch.cat([x_t, latents_cond], -1)
 * (src/core/infer.py:341) and, when augment != 0, the latent-noise augmentation
 * latent_blur = A*latent + B*aug_noise (src/core/generation_phases.py:685-693 -> schedules/base.py:82-87; A = 1 - t/T,
 * B = t/T after timestep_transform, fp32 in the reference, rounded to bf16 once).  augment == 0: aug_noise may be NULL.
 */
int svr2_sr_input_fwd(const void* noise, const void* latent, const void* aug_noise, float A, float B, int augment,
                      void* out, int64_t n_rows, int channels, void* stream);

/*
 * Endpoint of the Euler sampler for prediction type v_lerp, bf16 sampler state (every op rounds to bf16):
 * out = (x_t - B*pred) / (A + B).   Replaces EulerSampler.get_endpoint (src/common/diffusion/samplers/base.py:88-98) ->
 * Schedule.convert_from_pred (src/common/diffusion/schedules/base.py:107-109).  One-step schedule: A = 0, B = 1.
 * out may alias x_t or pred.
 */
int svr2_v_lerp_endpoint_fwd(const void* x_t, const void* pred, float A, float B, void* out, int64_t n, void* stream);

/*
 * GPU post-processing of a decoded batch: the pipeline's default colour correction and the output conversion.
 *
 * svr2_wavelet_level_fwd: one level of wavelet_decomposition (src/utils/color_fix.py:122-184) —
 *   low = wavelet_blur(src, radius): 3x3 kernel [1 2 1]^T [1 2 1] / 16 with dilation `radius`, replicate padding (a grouped
 *   conv2d in the reference; same tap order and fp32 accumulation as ATen's depthwise kernel);
 *   high = (high + src) - low, two separately rounded ops (`high_freq.add_(image).sub_(low_freq)`); `first` != 0: high starts at 0.
 *   src is indexed [T, C, H, W] through four ELEMENT strides (the pipeline hands a permuted view of its [T, H, W, C] video);
 *   low / high are planar contiguous [T, C, H, W]; high may be NULL (the style image only needs its low band).
 *   The caller passes radius = min(2^level, max(1, min(H, W) / 8)) (color_fix.py:136-139).
 * svr2_colorfix_finish_fwd: wavelet_reconstruction's combine + the batch normalisation + the writer's conversion —
 *   v = clamp(high + low_style, -1, 1) * 0.5 + 0.5  (color_fix.py:240-246, generation_phases.py:1337) -> out [T, H, W, C] (same
 *   dtype, may be NULL) and / or out_u8 [T, H, W, C] = (uint8)(v * 255) truncated like numpy's astype (inference_cli.py:691,734),
 *   with R and B swapped when bgr != 0 (cv2 order, inference_cli.py:693-695).
 * dtype: SVR2_DTYPE_F32 or SVR2_DTYPE_BF16 (the pipeline post-processes in its compute dtype, generation_phases.py:1229-1237);
 * bf16 rounds where the reference's bf16 tensors round.
 */
#define SVR2_DTYPE_F32  0
#define SVR2_DTYPE_BF16 1
#define SVR2_DTYPE_F16  2
int svr2_wavelet_level_fwd(const void* src, int64_t stride_t, int64_t stride_c, int64_t stride_y, int64_t stride_x, void* low,
                           void* high, int first, int T, int C, int H, int W, int radius, int dtype, void* stream);
int svr2_colorfix_finish_fwd(const void* high, const void* low_style, void* out, void* out_u8, int bgr, int T, int C, int H, int W,
                             int dtype, void* stream);

/*
 * The two colour-space conversions of lab_color_transfer (src/utils/color_fix.py:249-470, the CLI's default colour correction);
 * the histogram matching between them (color_fix.py:473-513: global sort + scatter per channel) stays on the host library's sort.
 * svr2_rgb_to_lab_fwd: rgb [T, 3, H, W] planar in [-1, 1] -> lab [T, 3, H, W] fp32 (_rgb_to_lab_batch after `.add_(1).mul_(0.5)
 *   .clamp_(0, 1)`); with `add` != NULL the input is clamp(rgb + add, -1, 1) — wavelet_reconstruction's combine, the first step of
 *   lab_color_transfer, fed with the high / low bands of svr2_wavelet_level_fwd.
 * svr2_lab_to_rgb_finish_fwd: lab -> rgb (_lab_to_rgb_batch) `.mul_(2).sub_(1)`, cast to the pipeline dtype, then the batch
 *   normalisation and writer conversion of svr2_colorfix_finish_fwd -> out [T, H, W, 3] / out_u8.
 * fp32 ops rounded one by one like the reference's tensor ops; the 3x3 colour matrices are summed in index order (the
 * reference's matmul order is the library's), so parity is within a few fp32 ulps, not bit-exact.
 */
int svr2_rgb_to_lab_fwd(const void* rgb, const void* add, float* lab, int T, int H, int W, int dtype, void* stream);
int svr2_lab_to_rgb_finish_fwd(const float* lab, void* out, void* out_u8, int bgr, int T, int H, int W, int dtype, void* stream);

/*
 * SURVEY §8f row N3 (first slice): the video VAE's causal 3-D convolution, stride 1 — InflatedCausalConv3d
 * (src/models/video_vae_v3/modules/causal_inflation_lib.py:61-262) as it is used by ResnetBlock3D.conv1 / conv2
 * (src/models/video_vae_v3/modules/attn_video_vae.py:263-280) and the up-sampler's convolutions.  Implicit GEMM on tcgen05:
 * no im2col buffer, the input is read by 4-D TMA boxes at the tap's offset (zero fill = the spatial zero padding).
 *   x [T, H, W, c_in], y / resid [T, H, W, c_out]: channels-last (torch.channels_last_3d of the reference's [1, C, T, H, W]);
 *   weight_packed [c_out, kt, ks, ks, c_in] from svr2_conv3d_pack_weight (nn.Conv3d's [c_out, c_in, kt, kh, kw], packed once);
 *   head: NULL = first temporal slice, the first frame is repeated kt - 1 times in front (extend_head, :422-437, basic_forward
 *     :228); else [kt - 1, H, W, c_in] = the last kt - 1 input frames of the previous slice (`self.memory`, :226-227, :254-272);
 *   bias [c_out] or NULL; resid or NULL: y = round(round(conv + bias) + resid) — ResnetBlock3D's `input_tensor + hidden_states`.
 *   kt in {1, 3}, ks in {1, 3}, c_in % 64 == 0, c_out % 128 == 0, dtype SVR2_DTYPE_F16 (the VAE's) or SVR2_DTYPE_BF16.
 */
int svr2_conv3d_pack_weight(const void* weight, void* packed, int c_out, int c_in, int kt, int kh, int kw, void* stream);
int svr2_causal_conv3d_fwd(const void* x, const void* head, const void* weight_packed, const void* bias, const void* resid,
                           void* y, void* gn_partials, int T, int H, int W, int c_in, int c_out, int kt, int ks, int dtype, void* stream);
/* gn_partials (or NULL): the epilogue also gathers the statistics of y that the VAE's next op, a per-frame GroupNorm over 32 groups
 * (ResnetBlock3D.norm2 / the next block's norm1), needs — T x svr2_conv3d_gn_slots() x 32 x 3 floats (count, mean, M2), zeroed and
 * filled by the call; hand them to svr2_groupnorm_silu_fwd (have_partials = svr2_conv3d_gn_slots()), whose statistics pass is
 * then skipped.  c_out in {128, 256, 512}. */
int svr2_conv3d_gn_slots(void);
/*
 * The encoder's down-sampling convolution: Downsample3D (attn_video_vae.py:176-246) = zero rows / columns after the last one
 * (`pad (0, 1, 0, 1)`, :239-241; pad_hw = 1 — or 0 if the caller has padded already) + InflatedCausalConv3d with kernel (kt, 3, 3), stride
 * (kt == 3 ? 2 : 1, 2, 2), no spatial padding.  x [T, H, W, c_in] -> y [T', H', W', c_out], H' = (H + pad_hw - 3) / 2 + 1 (W' alike),
 * T' = (T + head - 3) / 2 + 1 for kt == 3, where head = 2 copies of the first frame (head == NULL, first temporal slice) or the
 * head_frames (= kernel_t - stride_t = 1) frames the previous slice left in `memory`.  Each tap's input is one TMA box with element
 * strides 2; the padding is the unit's zero fill.
 */
int svr2_causal_conv3d_down_fwd(const void* x, const void* head, int head_frames, const void* weight_packed, const void* bias, void* y,
                                int T, int H, int W, int c_in, int c_out, int kt, int pad_hw, int dtype, void* stream);

/*
 * The VAE's per-frame GroupNorm (+ SiLU): causal_norm_wrapper folds time into the batch before nn.GroupNorm
 * (causal_inflation_lib.py:354-409), ResnetBlock3D applies SiLU to the result (attn_video_vae.py:316-322, 348-356).
 *   x, y [T, HW, C] channels-last (y may alias x), weight / bias [C] or NULL, statistics per (frame, group) in fp32, output rounded
 *   after the affine and again after the activation like the two ATen ops; silu = 0 skips the activation.
 *   workspace: svr2_groupnorm_workspace_bytes(T, groups) bytes of device memory (the statistics partials; no allocation inside),
 *   have_partials = 0; or the gn_partials a convolution gathered for this tensor, have_partials = svr2_conv3d_gn_slots().
 *   C / groups in {2, 4, 8, 16}, C = 8 x a power of two; dtype SVR2_DTYPE_F16 or SVR2_DTYPE_BF16.
 */
#define SVR2_GN_CHUNKS 128
int64_t svr2_groupnorm_workspace_bytes(int T, int groups);
int svr2_groupnorm_silu_fwd(const void* x, const void* weight, const void* bias, void* y, void* workspace, int have_partials, int T,
                            int64_t HW, int C, int groups, float eps, int silu, int dtype, void* stream);

/*
 * Layout change between the reference's NCDHW tensors and this library's channels-last VAE modules: dst[b][j][i] = src[b][i][j] for
 * 2-byte elements, i < rows, j < cols, with row pitches ld_src >= cols, ld_dst >= rows (elements) and batch strides.  [C, T*H*W] ->
 * [T*H*W, C]: rows = C, cols = T*H*W; the reverse with a channel slice (conv_out: 3 of 128): rows = T*H*W, cols = 3, ld_src = 128.
 * No counterpart in the reference (it stays NCDHW); replaces ATen's strided `.contiguous()` at the module boundaries.
 */
int svr2_transpose_2b(const void* src, void* dst, int64_t rows, int64_t cols, int64_t ld_src, int64_t ld_dst, int batch,
                      int64_t batch_stride_src, int64_t batch_stride_dst, void* stream);
/*
 * The up-samplers' pixel (and time) shuffle, Upsample3D.forward (attn_video_vae.py:134-153): einops
 * "b (x y z c) f h w -> b c (f z) (h x) (w y)" after the 1 x 1 x 1 expansion, and remove_head (the duplicated second frame of the
 * first temporal slice, causal_inflation_lib.py:412-419) — channels-last: src [T, H, W, spatial^2 * temporal * C] ->
 * dst [T * temporal - drop_second_frame, H * spatial, W * spatial, C], 2-byte elements, C % 8 == 0.
 */
int svr2_pixel_shuffle3d_fwd(const void* src, void* dst, int T, int H, int W, int C, int spatial, int temporal, int drop_second_frame,
                             void* stream);

/*
 * In-place softmax over the first n_valid columns of every row of a [rows, pitch] matrix of 2-byte elements; columns n_valid .. pitch - 1
 * become zeros.  The step between the two GEMMs (svr2_causal_conv3d_fwd as 1 x 1 x 1) of the VAE mid block's attention layer —
 * UNetMidBlock3D's diffusers `Attention` (attn_video_vae.py:563-668): one 512-channel head over the H x W positions of a frame, torch
 * SDPA in the reference.  pitch % 8 == 0.
 */
int svr2_row_softmax_fwd(void* x, int64_t rows, int n_valid, int pitch, int dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SEEDVR2_B200_H */
