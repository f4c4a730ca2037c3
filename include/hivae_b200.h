/*
 * hivae_b200 — C ABI of the B200 (sm_100a) hot path of Hi-VAE.
 *
 * Drop-in boundary: the reference (HuaizeLiu/Hi-VAE) has no FFI
 * layer of its own — its hot path is torch.nn.Module code that resolves to
 * cuBLASLt / SDPA / cuFFT / ATen library kernels.  Every entry point below names
 * the reference call site (path:line under /root/reference) whose device work it
 * replaces.  The Python host package `hivae_b200/` binds these symbols with
 * ctypes and mirrors the reference's nn.Module API.
 *
 * Conventions
 *   - plain pointers + sizes, no torch types; all pointers are DEVICE pointers
 *     unless the name ends in `_host`;
 *   - every function returns 0 on success, a negative HV_E* code on failure;
 *     `hv_last_error()` returns a thread-local message;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - no function allocates device memory; the caller owns inputs and outputs;
 *   - dtype codes: HV_F32 = 0, HV_BF16 = 1.
 *
 * Row addressing ("frames")
 *   Token streams live in an fp32 residual buffer laid out as `frames` frames of
 *   `frame_rows` rows x D columns.  A kernel works on the row range
 *   [row_off, row_off+L) of every frame; the "dense" matrix it reads/writes has
 *   frames*L rows, dense row r <-> (frame f = r / L, i = r % L) <-> residual row
 *   f*frame_rows + row_off + i.  Rows whose absolute in-frame index lies in
 *   [img_start, img_start+img_len) are "image" rows, all others "motion" rows
 *   (AMDTransformerBlock's two modulation sets, model/modules.py:1065-1109).
 */
#ifndef HIVAE_B200_H
#define HIVAE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HV_F32 0
#define HV_BF16 1

#define HV_OK 0
#define HV_EINVAL (-1)   /* bad argument / unsupported shape */
#define HV_ECUDA (-2)    /* CUDA runtime or driver error     */
#define HV_EARCH (-3)    /* device is not sm_100             */

const char* hv_last_error(void);
int hv_version(void);
/* 0 if the current device is compute capability 10.x, HV_EARCH otherwise. */
int hv_check_device(void);
/* number of kernels launched by this library since process start (bench `gpu_launches`). */
int64_t hv_launch_count(void);

/* ------------------------------------------------------------------ wavelet
 * model/wavelet.py:13-27 (dwt_init): one level of 2-D Haar analysis over
 * `planes` = B*C images of H x W (both even).  Outputs are planes x H/2 x W/2. */
int hv_dwt2_haar(const void* x, void* ll, void* hl, void* lh, void* hh,
                 int64_t planes, int H, int W, int dtype, void* stream);
/* Fused multi-level analysis: `levels` (1..3) applications of dwt_init to the running LL
 * (the caller-side recursion of model/frenquence_wavelet.py:86-89) in ONE pass over x.
 * hl/lh/hh: host arrays of `levels` device pointers (level k bands: planes x H/2^(k+1) x W/2^(k+1));
 * ll: LL of the last level.  Bit-identical to `levels` calls of hv_dwt2_haar in both dtypes.
 * Needs H % 2^levels == 0, W % 8 == 0, 16-byte aligned x. */
int hv_dwt2_haar_pyramid(const void* x, int levels, void* const* hl, void* const* lh, void* const* hh, void* ll,
                         int64_t planes, int H, int W, int dtype, void* stream);
/* Fused multi-level synthesis, the inverse of hv_dwt2_haar_pyramid: `levels` applications of
 * iwt_init starting from the coarsest LL.  H x W is the full-resolution OUTPUT size, out is fp32.
 * Bit-identical to the per-level path (coarsest level in the band dtype, later levels in fp32). */
int hv_idwt2_haar_pyramid(const void* ll, int levels, const void* const* hl, const void* const* lh, const void* const* hh,
                          float* out, int64_t planes, int H, int W, int dtype, void* stream);
/* model/wavelet.py:31-48 (iwt_init): inverse; the four sub-bands are the four
 * dim-0 quarters of the reference's input.  Output is ALWAYS fp32 (wavelet.py:40-41).
 * dtype_in = the dtype of all four bands, or dtype_ll | (dtype_bands + 1) << 8 when LL differs from the
 * detail bands (second level of a bf16 pyramid: torch.cat promotes the mixed list to fp32). */
int hv_idwt2_haar(const void* ll, const void* hl, const void* lh, const void* hh, float* out,
                  int64_t planes, int h, int w, int dtype_in, void* stream);

/* ------------------------------------------------------------------ frequency split
 * model/frequency_utils.py:84-108 (freq_3d_filter) with the Gaussian mask of
 * frequency_utils.py:45-81.  Because the mask is real, even and separable (even
 * T,H,W), ifftn(fftn(x)*LPF).real equals three 1-D circular convolutions;
 * ct/ch/cw are the T x T, H x H, W x W circulant matrices (fp32, row-major) the
 * host derives from the mask.  x, lo, hi: vols x T x H x W fp32; hi = x - lo.
 * Requires T*H*W*4 bytes <= 200 KB (one volume per CTA in shared memory).
 * T in {8,16,32} with H = W in {16,32} and 16-byte aligned buffers run the three
 * passes on the tensor cores (3xTF32 split, ~1e-6 relative; the reference's FFT
 * path agrees to 1e-5); everything else, and 32^3 launches too small to fill the
 * SMs, runs the fp32 SIMT kernels.  Environment HV_FREQ_TC=0 / =1 forces the SIMT
 * / the tensor kernel (A/B measurements, tools/bench_freq.py). */
int hv_freq_split3d(const float* x, const float* ct, const float* ch, const float* cw,
                    float* lo, float* hi, int64_t vols, int T, int H, int W, void* stream);

/* ------------------------------------------------------------------ row kernels
 * LayerNorm (+ optional adaLN modulate) of residual rows -> dense activation.
 *   y = (LN(x) * ln_w + ln_b) * (1 + scale) + shift
 * Replaces nn.LayerNorm / AMDLayerNormZero / AMDLayerNormZero_OneVariable
 * (model/modules.py:697,710,881-928) and the torch.cat of the two streams
 * (modules.py:1088,1103).  `ada` (may be NULL) is the fp32 table
 * Linear(SiLU(temb)) with row stride ada_ld; the conditioning row of frame f is
 * (f / cond_div) * cond_mul; shift/scale column offsets are given per segment. */
int hv_ln_modulate(const float* res, int64_t frames, int64_t frame_rows, int row_off, int L, int D,
                   const float* ln_w, const float* ln_b, float eps,
                   const float* ada, int64_t ada_ld, int shift_col_mot, int scale_col_mot,
                   int shift_col_img, int scale_col_img, int img_start, int img_len,
                   int cond_div, int cond_mul, void* out, int out_dtype, void* stream);

/* the same, for the training forward: also writes the fp32 rows entering the norm to `save` (dense frames*L x D) */
int hv_ln_modulate_save(const float* res, int64_t frames, int64_t frame_rows, int row_off, int L, int D,
                        const float* ln_w, const float* ln_b, float eps,
                        const float* ada, int64_t ada_ld, int shift_col_mot, int scale_col_mot,
                        int shift_col_img, int scale_col_img, int img_start, int img_len,
                        int cond_div, int cond_mul, void* out, int out_dtype, float* save, void* stream);

/* PatchEmbed (model/modules.py:858-879) = Conv2d(k=s=p) + bias, fused with the
 * 2-D sincos and temporal sincos adds of transformer.py:1985-1988 / 130-141 /
 * 506-514.  Input channels come from two NCHW tensors (in1: C1 channels, in2: C2
 * channels, in2 may be NULL) — the torch.cat((zi,zt),1) of model_AMD.py:2178.
 * Token (f, s) is written to row f*frame_rows + row_off + s of `out` (fp32 or
 * bf16, row stride D).  pos: S x D fp32 or NULL; tpos: Tmod x D fp32 or NULL,
 * indexed by f % Tmod.  w: D x (C1+C2) x p x p fp32; bias: D. */
int hv_patch_embed(const float* in1, int C1, const float* in2, int C2, int64_t frames, int H, int W, int p,
                   const float* w, const float* bias, const float* pos, const float* tpos, int Tmod,
                   void* out, int out_dtype, int64_t frame_rows, int row_off, int D, void* stream);

/* DiT head (model/transformer.py:2032-2041): norm_final -> AdaLayerNorm(chunk_dim=1,
 * order shift,scale; modules.py:1001-1012) -> Linear(D -> C*p*p) -> unpatchify with
 * (c,p,q) feature order -> optional Euler update zt += dt*pre (model_AMD.py:2156).
 * pre / zt: frames x C x H x W fp32 (zt may be NULL). */
int hv_final_head(const float* res, int64_t frames, int64_t frame_rows, int row_off, int S, int D,
                  const float* n1_w, const float* n1_b, const float* n2_w, const float* n2_b, float eps,
                  const float* ada, int64_t ada_ld, int shift_col, int scale_col,
                  const float* w, const float* bias, int C, int p, int Wp,
                  float* pre, float* zt, float dt, void* stream);

/* Timesteps(dim, flip_sin_to_cos=True, freq_shift=0) (diffusers; call site
 * model/transformer.py:1939): out[r] = [cos(t*f_i) | sin(t*f_i)], fp32 or bf16. */
int hv_timestep_embed(const float* t, int64_t rows, int dim, void* out, int out_dtype, void* stream);

/* elementwise helpers: y = silu(x) (adaLN / TimestepEmbedding activations),
 * dtype conversion, y = a*x + y, strided row gather/scatter between a dense
 * matrix and residual frames (builds the token streams of transformer.py:1971-1979). */
int hv_silu(const void* x, int in_dtype, void* y, int out_dtype, int64_t n, void* stream);
int hv_convert(const void* x, int in_dtype, void* y, int out_dtype, int64_t n, void* stream);
int hv_axpy(float a, const float* x, float* y, int64_t n, void* stream);
/* res[f*frame_rows + row_off + i, :] = src[f*src_stride_f + i*src_stride_i, :] + add[i, :]
 * (strides in rows of D floats; src_stride_f = 0 broadcasts one token set; add may be NULL, L x D). */
int hv_rows_scatter(const float* src, int64_t src_stride_f, int64_t src_stride_i, const float* add,
                    float* res, int64_t frames, int64_t frame_rows, int row_off, int L, int D, void* stream);
/* dense fp32 out[(f*L+i), :] = res[f*frame_rows + row_off + i, :] */
/* ---- building blocks of the backward pass (autograd of nn.Linear / LayerNorm / GELU under `accelerator.backward`,
 * train_amd.py:489; forward definitions model/modules.py:651-1188).  The gradient GEMMs themselves go through
 * hv_linear_bf16 with transposed operands:  dX = dY W  (A = dY, W := W^T),  dW = dY^T X  (A = dY^T, W := X^T). */
/* out[c][r] = in[r][c]; dtype HV_F32 / HV_BF16; ld in elements */
int hv_transpose2d(const void* in, void* out, int64_t rows, int64_t cols, int64_t ld_in, int64_t ld_out, int dtype, void* stream);
/* out = in^T and colsum[c] = sum_r in[r][c] in one pass (dY^T + bias gradient of nn.Linear's backward); bf16, sizes and
 * strides multiples of 8, 16-byte aligned bases; workspace = ceil(rows / 64) * cols floats */
int hv_transpose2d_colsum(const void* in, void* out, int64_t rows, int64_t cols, int64_t ld_in, int64_t ld_out, int dtype,
                          float* workspace, float* colsum, void* stream);
/* bias gradient: out[n] = sum_m x[m][n]; workspace = hv_colsum_slices() * N floats */
int hv_colsum_slices(void);
int hv_colsum(const void* x, int64_t M, int N, int64_t ld, int dtype, float* workspace, float* out, void* stream);
/* hv_colsum into up to three destinations (columns [0, n0) -> out0, [n0, n0 + n1) -> out1, the rest -> out2; n0 = N for a single
 * one), added to what is there when accumulate != 0: bias gradients straight into their .grad slots (the q / k / v biases of a
 * packed projection are separate tensors), as loss.backward() accumulates at train_amd.py:489 */
int hv_colsum_acc(const void* x, int64_t M, int N, int64_t ld, int dtype, float* workspace, float* out0, int n0, float* out1,
                  int n1, float* out2, int accumulate, void* stream);
/* dx = dy * d/dx gelu_tanh(pre)  (diffusers GELU(approximate="tanh") inside FeedForward, modules.py:712-719) */
int hv_gelu_tanh_bwd(const void* pre, const void* dy, void* dx, int64_t n, int dtype_in, int dtype_out, void* stream);
/* nn.LayerNorm backward over dense fp32 rows (D % 128 == 0, D <= 1024): dx written (or accumulated), dw / db optional.
 * workspace = (hv_layernorm_bwd_ctas() + 1) * 2 * D floats */
int hv_layernorm_bwd_ctas(void);
int hv_layernorm_bwd(const float* x, const void* dy, int dy_dtype, const float* w, float eps, int64_t rows, int D,
                     float* dx, int accumulate, float* workspace, float* dw, float* db, void* stream);

/* backward of the per-head LayerNorm(64) applied to q / k (diffusers Attention(qk_norm="layer_norm"), eps 1e-6,
 * modules.py:699-707): x, dy, dx are (tokens, heads*64) views, one dtype; dw / db (64 floats) are summed over tokens and
 * heads.  workspace = (hv_layernorm_bwd_ctas() + 1) * 128 floats */
int hv_headnorm64_bwd(const void* x, int64_t ldx, const void* dy, int64_t lddy, const float* w, float eps, int64_t tokens,
                      int heads, void* dx, int64_t lddx, int dtype, float* workspace, float* dw, float* db, void* stream);
/* hv_headnorm64_bwd with dw / db added to their destinations when accumulate != 0 */
int hv_headnorm64_bwd_acc(const void* x, int64_t ldx, const void* dy, int64_t lddy, const float* w, float eps, int64_t tokens,
                      int heads, void* dx, int64_t lddx, int dtype, float* workspace, float* dw, float* db, int accumulate, void* stream);

/* ---- backward of the adaLN-Zero blocks (AMDTransformerBlock / BasicDiTBlock, model/modules.py:1065-1109, 1160-1188) and of
 * the small layers around them; csrc/train_ops.cu.  Same row addressing as the forward ("frames" above).  Every reduction
 * is two-stage in a fixed order (bit-reproducible).  hv_train_chunk_rows() = rows per CTA chunk: a frame's range is cut into
 * G <= L / chunk + 3 chunks, which sizes the workspaces below. */
int hv_train_chunk_rows(void);
/* gate backward of `hidden = hidden + gate * branch` (modules.py:1094-1107): dY[dense row] = gate * dres[row] in `dtype`
 * (fp32 / bf16), dtable[cond row, gate column of the segment] += sum_rows dres * y (y = the saved branch output, dense,
 * `dtype`).  gate == NULL: plain residual branch, dY = dres rows (gather + convert), y / dtable / workspace unused.
 * workspace >= frames * G * D floats. */
int hv_gate_bwd(const float* dres, int64_t frames, int64_t frame_rows, int row_off, int L, int D, const void* y,
                const float* gate, int64_t gate_ld, int gate_col_mot, int gate_col_img, int img_start, int img_len,
                int cond_div, int cond_mul, void* dY, int dtype, float* dtable, float* workspace, int64_t ws_floats,
                void* stream);
/* backward of hv_ln_modulate: x = the saved fp32 input rows (dense, frames*L x D), dy dense (fp32 / bf16).  The input
 * gradient is ACCUMULATED into dres rows [row_off, row_off+L) of every frame; dtable (same layout as `ada`) accumulates
 * dshift / dscale per conditioning row and segment; dw / db (D floats, may be NULL) are accumulated (+=).
 * ada == NULL: plain nn.LayerNorm backward.  workspace >= (frames * G + groups) * 2 * D floats, groups = ceil(frames /
 * cond_div) (cond_div = 8 when ada == NULL). */
int hv_ln_modulate_bwd(const float* x, const void* dy, int dy_dtype, const float* ln_w, const float* ln_b, float eps,
                       float* dres, int64_t frames, int64_t frame_rows, int row_off, int L, int D, const float* ada,
                       int64_t ada_ld, int shift_col_mot, int scale_col_mot, int shift_col_img, int scale_col_img,
                       int img_start, int img_len, int cond_div, int cond_mul, float* dtable, float* dw, float* db,
                       float* workspace, int64_t ws_floats, void* stream);
/* training forward of the gated residual when the branch output must be kept: res[row] += gate * y[dense row]
 * (gate == NULL: plain add); the inference path fuses this into the GEMM epilogue instead */
int hv_gate_residual(float* res, int64_t frames, int64_t frame_rows, int row_off, int L, int D, const void* y, int dtype,
                     const float* gate, int64_t gate_ld, int gate_col_mot, int gate_col_img, int img_start, int img_len,
                     int cond_div, int cond_mul, void* stream);
/* dx = dy * silu'(x) (x fp32): TimestepEmbedding / adaLN activations (transformer.py:1940-1941, modules.py:894) */
int hv_silu_bwd(const float* x, const void* dy, int dy_dtype, void* dx, int dx_dtype, int64_t n, void* stream);
/* dpre = gout[0] * 2/n * (pre - gt): backward of hv_mse_loss (gout: device scalar or NULL = 1) */
int hv_mse_bwd(const float* pre, const float* gt, int64_t n, const float* gout, float* dpre, void* stream);
/* out[o][:] (+)= sum_{a<A} sum_{b<B} src[base + o*stride_out + a*stride_a + b*stride_b][:]  (rows of D floats, fixed order):
 * gradients of broadcast token rows (motion_token, source_token / target_token, transformer.py:122-123, 1971-1979) */
int hv_rows_reduce(const float* src, int D, int64_t n_out, int64_t base, int64_t stride_out, int64_t A, int64_t stride_a,
                   int64_t B, int64_t stride_b, float* out, int accumulate, void* stream);
/* random_masking's gather (transformer.py:181-183, 575-577): out[f, j, :] = in[f, idx[f*keep + j], :] (frame strides in
 * rows); scatter != 0 is its backward out[f, idx[..], :] = in[f, j, :] into a pre-zeroed buffer */
int hv_rows_index(const float* in, float* out, const int* idx, int64_t frames, int S, int keep, int D, int64_t in_frame_rows,
                  int64_t out_frame_rows, int scatter, void* stream);
/* camera_down backward (model/model_AMD.py:1788-1793): dx written (may be NULL), dw / db accumulated (may be NULL) */
int hv_conv3x3_bwd(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db, int64_t B, int Cin,
                   int Cout, int H, int W, int stride, void* stream);
int hv_maxpool2_bwd(const float* x, const float* dy, float* dx, int64_t planes, int H, int W, void* stream);

/* Diagnostics: with HV_ATTN_PROF=1 in the environment the tcgen05 attention kernel accumulates per-phase cycle
 * counters (148 CTAs x 2 key halves x 8 u64: wait S, max pass, max exchange, O epilogue, exp pass, publish P, full
 * tiles, tail tiles); this copies them to `host` and clears them.  Returns 0 when profiling is off. */
int hv_attn_prof_read(unsigned long long* host);
/* HV_ATTN_TIMELINE=1: clock64 stamps of CTA 0 of the last attn_tc2 launches (64 tiles x 16 events) -> host[1024]; 0 when off */
int hv_attn_timeline_read(unsigned long long* host);
/* the same for the tcgen05 attention BACKWARD kernels (HV_ATTN_BWD_PROF=1): 2 kernels (dq, dkv) x 148 CTAs x 16 u64 */
int hv_attn_bwd_prof_read(unsigned long long* host);

/* tokens (N, h*w, p*p*c) <-> image (N, c, h*p, w*p).  order 0 = (p,q,c) features: AMDModel*.unpatchify
 * (model_AMD.py:2816-2830), MaskedAutoencoderViT.patchify/unpatchify (model_MAE.py:116-145); order 1 = (c,p,q): the DiT
 * heads (transformer.py:2038-2041).  inverse != 0: image -> tokens (patchify). */
int hv_unpatchify(const float* in, float* out, int64_t N, int c, int h, int w, int p, int order, int inverse, void* stream);

int hv_rows_gather(const float* res, int64_t frames, int64_t frame_rows, int row_off, int L, int D,
                   float* out, void* stream);

/* camera_down (model/model_AMD.py:1788-1793): direct 3x3 convolution (NCHW fp32,
 * zero padding 1, stride 1 or 2) and 2x2 max-pool. */
int hv_conv3x3(const float* x, const float* w, const float* bias, float* y,
               int64_t B, int Cin, int Cout, int H, int W, int stride, void* stream);
int hv_maxpool2(const float* x, float* y, int64_t planes, int H, int W, void* stream);

/* RoPE of model/attention.py:39-62 applied in place to q and k rows:
 * pairs (2i,2i+1) of every head rotated by angle (start_pos+pos) * base^(-2i/head_dim). */
int hv_rope(float* q, float* k, int64_t batch, int L, int heads, int head_dim, int64_t ld,
            int start_pos, float base, void* stream);

/* ------------------------------------------------------------------ GEMM + fused epilogues
 * out = epilogue(A[M,K] @ W[N,K]^T)   (nn.Linear layout, W row-major N x K).
 * Replaces the to_q/to_k/to_v/to_out Linears of diffusers Attention, FeedForward's
 * two Linears, the adaLN / time-embedding Linears and proj_out
 * (call sites model/modules.py:699-719,894,901; transformer.py:1940-1941,2036). */
typedef struct HvEpilogue {
    int mode;            /* HV_EPI_* */
    int out_dtype;       /* HV_F32 / HV_BF16 (modes STORE, GELU, QKNORM) */
    void* out;           /* dense output, row stride ldo elements */
    int64_t ldo;
    const float* bias;   /* N or NULL */
    /* HV_EPI_RESIDUAL: res[row(r), :] += gate * (acc + bias) */
    float* res;
    int L;
    int64_t frame_rows;
    int row_off;
    const float* gate;   /* adaLN table or NULL (plain residual add) */
    int64_t gate_ld;
    int gate_col_mot, gate_col_img;
    int img_start, img_len;
    int cond_div, cond_mul;
    /* HV_EPI_QKNORM: per-head (64-wide) LayerNorm on q columns [0,q_cols) and k columns
     * [q_cols, q_cols+k_cols); remaining columns (v) stored as-is.
     * (diffusers Attention norm_q/norm_k, eps 1e-6, model/modules.py:703-704) */
    const float *qn_w, *qn_b, *kn_w, *kn_b;
    int q_cols, k_cols;
    float qk_eps;
    /* HV_EPI_RESIDUAL on the tcgen05 path only: split the K dimension over k_split work items per output tile, each
     * reduce-adding its partial product into res (weight gradients: few output tiles, very long K).  0 / 1 = off.
     * Needs bias == gate == NULL; fp32 partials are added in arrival order. */
    int k_split;
    /* modes GELU / QKNORM, optional second output (training forward): the pre-activation matrix acc + bias (before the GELU /
     * before the per-head LayerNorm) in the same dtype as `out`, row stride ldo2; NULL = off.  One GEMM then yields both the
     * activations the next layer consumes and what the backward needs (GELU', QK-norm backward). */
    void* out2;
    int64_t ldo2;
} HvEpilogue;

#define HV_EPI_STORE 0
#define HV_EPI_GELU 1      /* gelu(tanh) — FeedForward "gelu-approximate" */
#define HV_EPI_RESIDUAL 2
#define HV_EPI_QKNORM 3

/* fp32 SIMT GEMM (parity mode and skinny/tiny-K linears). A: fp32 or bf16, W fp32. */
int hv_linear_f32(const void* A, int a_dtype, int64_t lda, const float* W, int64_t ldw,
                  int64_t M, int N, int K, const HvEpilogue* epi, void* stream);
/* bf16 tcgen05/TMEM GEMM fed by TMA (fp32 accumulate). A, W bf16; K % 8 == 0,
 * lda/ldw % 8 == 0, 16-byte aligned bases. */
int hv_linear_bf16(const void* A, int64_t lda, const void* W, int64_t ldw,
                   int64_t M, int N, int K, const HvEpilogue* epi, void* stream);

/* Weight-gradient form of the same kernel: out[M x N] (+)= At^T Bt with At (K x M) and Bt (K x N) row-major bf16 — dW = dY^T X
 * of nn.Linear's backward with dY and X used as they lie in memory (MN-major tcgen05 operands, no transposed copies).
 * K (tokens) arbitrary; M, N, ldat, ldbt % 8 == 0, 16-byte aligned bases.  Epilogue: HV_EPI_STORE, or HV_EPI_RESIDUAL
 * without bias / gate (accumulates into fp32, `k_split` allowed). */
int hv_linear_bf16_tn(const void* At, int64_t ldat, const void* Bt, int64_t ldbt, int64_t M, int N, int64_t K,
                      const HvEpilogue* epi, void* stream);

/* ------------------------------------------------------------------ attention
 * softmax(Q K^T * scale) V per (sequence, head), head_dim = hd, no dropout
 * (F.scaled_dot_product_attention inside diffusers AttnProcessor2_0; call sites
 * model/modules.py:699-707,794-803,1043-1051,1138-1146; model/attention.py:105-112).
 * Row of token i of sequence s:  (s / sdiv)*sa + (s % sdiv)*sb + i*sc  (separately
 * for the q/out side and the k/v side).  q/k/v/out point at column 0 of head 0;
 * head h starts at column h*hd.  mask: optional Lq x Lkv fp32, entries == 0 are
 * filled with -1e9 before the softmax (attention.py:108-109).  dtype applies to
 * q,k,v,out. */
typedef struct HvSeqMap { int sdiv; int64_t sa, sb, sc; } HvSeqMap;
int hv_attention(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                 void* out, int64_t ldo, int dtype, int64_t seqs, int heads, int hd, int Lq, int Lkv,
                 const HvSeqMap* qmap, const HvSeqMap* kvmap, float scale, const float* mask, void* stream);

/* Attention backward (bf16: warp-level tensor-core kernels, fp32 accumulation; fp32: SIMT kernels): gradients of out = softmax(q k^T * scale) v for head_dim 64, no
 * mask — autograd of F.scaled_dot_product_attention at model/modules.py:699-707, 1043-1051.  Same matrix / sequence-map
 * conventions as hv_attention; `stats` = seqs * heads * Lq * 2 floats of workspace (lse, delta per query row). */
int hv_attention_bwd(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                     const void* out, int64_t ldo, const void* dout, int64_t lddo,
                     void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv, int dtype,
                     int64_t seqs, int heads, int hd, int Lq, int Lkv, const HvSeqMap* qmap, const HvSeqMap* kvmap,
                     float scale, float* stats, void* stream);
/* Forward that hands its per-row log-sum-exp to the backward (the training forward of the blocks, model/modules.py:699-707,
 * 1043-1051 under autograd): hv_attention + `stats` = the workspace hv_attention_bwd takes.  Returns 1 when element 0 of every
 * (sequence, head, query row) pair of `stats` was written (tcgen05 kernel attn_tc2.cu), 0 when the kernel that ran does not
 * produce it (the output is complete either way), < 0 on error. */
int hv_attention_lse(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                     void* out, int64_t ldo, int dtype, int64_t seqs, int heads, int hd, int Lq, int Lkv,
                     const HvSeqMap* qmap, const HvSeqMap* kvmap, float scale, const float* mask, float* stats, void* stream);
/* hv_attention_bwd after hv_attention_lse returned 1 (have_lse = 1): the tcgen05 dq kernel skips its row-maximum pass and both
 * cross-warp exchanges; have_lse = 0 is hv_attention_bwd. */
int hv_attention_bwd_lse(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                     const void* out, int64_t ldo, const void* dout, int64_t lddo,
                     void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv, int dtype,
                     int64_t seqs, int heads, int hd, int Lq, int Lkv, const HvSeqMap* qmap, const HvSeqMap* kvmap,
                     float scale, float* stats, int have_lse, void* stream);


/* ------------------------------------------------------------------ train-step tail (train_amd.py:489-496)
 * Flat fp32 arenas (parameters, gradients, AdamW moments), no host synchronisation.
 * workspace: hv_reduce_workspace_doubles() fp64 values of device scratch. */
int hv_reduce_workspace_doubles(void);
/* out[0] = mean((pre-gt)^2): model/loss.py:13-14 (F.mse_loss), the diff_loss / rec_loss of model_AMD.py:1997-2000 */
int hv_mse_loss(const float* pre, const float* gt, int64_t n, double* workspace, float* out, void* stream);
/* hv_mse_loss without a caller-provided workspace and without ANY allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, serving loops).
 * The fp64 partial sums live in a static device array of 1024 doubles (8 KB per GPU that loads the library, until exit);
 * inputs that need more (n > hv_mse_loss_static_max_elements() = 1 048 576 on a B200) are rejected with HV_EINVAL.
 * Same result as hv_mse_loss, bit for bit.  All calls share the array: calls that overlap on the GPU race.  The first
 * call on a device looks the array's address up (cudaGetSymbolAddress): make it before a graph capture begins. */
int hv_mse_loss_static(const float* pre, const float* gt, int64_t n, float* out, void* stream);
int hv_mse_loss_static_max_elements(void);
/* out[0] = ||grad_scale*g||_2, out[1] = min(1, max_norm/(out[0]+1e-6)): torch.nn.utils.clip_grad_norm_
 * behind accelerator.clip_grad_norm_ (train_amd.py:491-492).  out: 2 floats on the device. */
int hv_grad_norm_clip_coef(const float* g, int64_t n, float grad_scale, float max_norm, double* workspace,
                           float* out, void* stream);
/* fused gradient clip + torch.optim.AdamW step `step` (1-based) (train_amd.py:255,494; defaults
 * betas (0.9, 0.999), eps 1e-8, weight_decay 0.01), same fp32 operation order as torch's single-tensor path.
 * g is read as grad_scale * clip_coef[0] * g (clip_coef may be NULL). 28 B/parameter of HBM traffic.
 * Hyper-parameters are doubles: torch forms 1-beta, 1-lr*wd, lr/bias_correction in Python doubles and rounds once. */
int hv_adamw_clip(float* p, const float* g, float* m, float* v, int64_t n, double lr, double beta1, double beta2,
                  double eps, double weight_decay, int64_t step, float grad_scale, const float* clip_coef, void* stream);
/* hv_adamw_clip with the step count and learning rate in device memory (state: 2 doubles {steps done, lr}; derived: 3 floats
 * of scratch): every launch advances the step by one, so a captured CUDA graph of the train step (train_amd.py:459-496)
 * can be replayed; same update as hv_adamw_clip */
int hv_adamw_clip_devstate(float* p, const float* g, float* m, float* v, int64_t n, double* state, float* derived, double beta1,
                           double beta2, double eps, double weight_decay, float grad_scale, const float* clip_coef, void* stream);
/* x *= coef[0] * extra: materialises the clip in the gradient arena (torch's observable in-place scaling) */
int hv_scale_by_device(float* x, int64_t n, const float* coef, float extra, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HIVAE_B200_H */
