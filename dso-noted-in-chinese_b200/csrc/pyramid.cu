// K1 -- image / gradient pyramid, replaces FrameHessian::makeImages
// (reference: FullSystem/HessianBlocks.cpp:137-211).
//
// One launch per frame, one pass over the input (pyr_fused_kernel): each CTA owns a 64x64 tile of the input image and
// the matching tiles of all coarser levels.  It brings the tile plus a halo of 2^(levels-1) pixels into shared memory --
// with one TMA tile load (cp.async.bulk.tensor.2d, out-of-image elements zero-filled by the hardware) when the image
// pitch and address allow, with plain loads otherwise -- cascades the exact 2x2 box sums 0.25f*(((a+b)+c)+d) (:179-182)
// through shared memory, so that level l+1 is bit-identical to the reference's level-by-level loop, and computes the
// gradients of EVERY level from the shared-memory tiles (the halo shrinks to one pixel at the coarsest level), writing
// (I,dx,dy,|grad|^2) as one float4 per pixel -- the device-resident layout every gather in the tracker / BA uses -- and, on
// request, the reference's 12-byte AoS + absSquaredGrad arrays for the host copy.  The halo pixels are re-read through L2
// (they are the neighbouring CTAs' tiles), DRAM sees the input once.  The reference's dx at x = 0 / w-1 reads the pixel of
// the neighbouring row (linear index +-1): those few pixels are rebuilt from the input by the exact cascade.
// The two-launch form (pyr_intensity_kernel + pyr_gradient_kernel) remains for DSO_B200_K1_TWO_PASS=1 (A/B timing).
// HBM-bound stencil: algorithmic bytes per frame = 4*w*h in + 16*sum_l(w_l*h_l) out.
#include "common.cuh"
#include <cuda.h>

#define TILE 64

struct IntensityArgs
{
	const float* src;
	int w, h, levels;
	float* inten[MAXL];   // inten[0] unused (the source is level 0)
};

__global__ void __launch_bounds__(256) pyr_intensity_kernel(IntensityArgs a)
{
	__shared__ float t0[TILE * TILE];
	__shared__ float t1[(TILE / 2) * (TILE / 2)];
	const int x0 = blockIdx.x * TILE, y0 = blockIdx.y * TILE;
	const int tid = threadIdx.x;
	const bool vec = ((a.w & 3) == 0) && ((reinterpret_cast<size_t>(a.src) & 15) == 0);
	if (vec)
	{
		// 64 rows x 16 float4
		for (int i = tid; i < TILE * TILE / 4; i += 256)
		{
			int r = i >> 4, c4 = (i & 15) << 2;
			int x = x0 + c4, y = y0 + r;
			float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
			if (y < a.h && x < a.w) v = __ldg(reinterpret_cast<const float4*>(a.src + (size_t)y * a.w + x));
			*reinterpret_cast<float4*>(&t0[r * TILE + c4]) = v;
		}
	}
	else
	{
		for (int i = tid; i < TILE * TILE; i += 256)
		{
			int r = i >> 6, c = i & 63;
			int x = x0 + c, y = y0 + r;
			t0[i] = (y < a.h && x < a.w) ? __ldg(a.src + (size_t)y * a.w + x) : 0.f;
		}
	}
	__syncthreads();
	float* cur = t0;
	float* nxt = t1;
	int ts = TILE;   // tile size at the current (source) level
	for (int l = 1; l < a.levels; l++)
	{
		const int td = ts >> 1;
		const int wl = a.w >> l, hl = a.h >> l;
		const int gx0 = x0 >> l, gy0 = y0 >> l;
		for (int i = tid; i < td * td; i += 256)
		{
			int yy = i / td, xx = i - yy * td;
			const float* p = cur + (2 * yy) * ts + 2 * xx;
			// HessianBlocks.cpp:179-182: 0.25f * (a + b + c + d), left to right
			float s = fmul(0.25f, fadd(fadd(fadd(p[0], p[1]), p[ts]), p[ts + 1]));
			nxt[yy * td + xx] = s;
			int gx = gx0 + xx, gy = gy0 + yy;
			if (gx < wl && gy < hl) a.inten[l][(size_t)gy * wl + gx] = s;
		}
		__syncthreads();
		float* t = cur; cur = nxt; nxt = t;
		ts = td;
	}
}

struct GradLevel
{
	const float* inten;
	float4* out;
	float* aos;      // may be null
	float* absg;     // may be null
	int w, h;
	int block_start; // first blockIdx.x of this level
	int vec;
};
struct GradArgs
{
	GradLevel lv[MAXL];
	int levels;
	const float* Blut;   // may be null
};

__device__ __forceinline__ void grad_px(float c, float l, float r, float u, float d, const float* Blut,
                                        bool interior, float& dx, float& dy, float& ab)
{
	if (!interior) { dx = dy = ab = 0.f; return; }   // rows 0 / h-1: never written by the reference
	dx = fmul(0.5f, fsub(r, l));     // :190
	dy = fmul(0.5f, fsub(d, u));     // :191
	if (!isfinite(dx)) dx = 0.f;
	if (!isfinite(dy)) dy = 0.f;
	ab = fadd(fmul(dx, dx), fmul(dy, dy));   // :201
	if (Blut)
	{
		// CalibHessian::getBGradOnly, HessianBlocks.h:402-410
		// int c = color + 0.5f; a non-finite / out-of-range value converts to INT_MIN on x86 (cvttss2si), i.e. clamps to 5
		const float cf = fadd(c, 0.5f);
		int ci = (cf >= -2147483648.f && cf < 2147483648.f) ? (int)cf : 5;
		ci = ci < 5 ? 5 : (ci > 250 ? 250 : ci);
		float gw = fsub(__ldg(Blut + ci + 1), __ldg(Blut + ci));
		ab = fmul(ab, fmul(gw, gw));   // :207
	}
}

__global__ void __launch_bounds__(256) pyr_gradient_kernel(GradArgs a)
{
	int l = 0;
#pragma unroll
	for (int k = 1; k < MAXL; k++)
		if (k < a.levels && (int)blockIdx.x >= a.lv[k].block_start) l = k;
	const GradLevel L = a.lv[l];
	const int n = L.w * L.h;
	const int wl = L.w;
	const int lo = wl, hi = wl * (L.h - 1);   // :188 gradient range
	const float* I = L.inten;
	// 1024 consecutive pixels per CTA; thread t takes pixels t, t+256, t+512, t+768 of them, so every load and every
	// 16-byte store instruction of a warp covers 32 consecutive pixels (128 B read, 512 B written, full sectors)
	const int start = (blockIdx.x - L.block_start) * 1024 + threadIdx.x;
	float c[4], lf[4], rt[4], u[4], d[4];
	bool interior[4], valid[4];
#pragma unroll
	for (int k = 0; k < 4; k++)
	{
		const int idx = start + 256 * k;
		valid[k] = idx < n;
		interior[k] = valid[k] && idx >= lo && idx < hi;
		c[k] = valid[k] ? __ldg(I + idx) : 0.f;
		lf[k] = rt[k] = u[k] = d[k] = 0.f;
		if (interior[k]) { lf[k] = __ldg(I + idx - 1); rt[k] = __ldg(I + idx + 1); u[k] = __ldg(I + idx - wl); d[k] = __ldg(I + idx + wl); }
	}
#pragma unroll
	for (int k = 0; k < 4; k++)
	{
		if (!valid[k]) continue;
		const int idx = start + 256 * k;
		float dx, dy, ab;
		grad_px(c[k], lf[k], rt[k], u[k], d[k], a.Blut, interior[k], dx, dy, ab);
		L.out[idx] = make_float4(c[k], dx, dy, ab);
		if (L.aos) { L.aos[3 * (size_t)idx] = c[k]; L.aos[3 * (size_t)idx + 1] = dx; L.aos[3 * (size_t)idx + 2] = dy; }
		if (L.absg) L.absg[idx] = ab;
	}
}

// ---- single-pass kernel ---------------------------------------------------------------------------------------
struct FusedArgs
{
	const float* src;
	int w, h, levels, H;      // H: halo of the level-0 tile (a multiple of 2^(levels-1), >= 4)
	float4* out[MAXL];
	float* aos[MAXL];         // may be null
	float* absg[MAXL];        // may be null
	const float* Blut;        // may be null
};

// gradients of one level's owned tile (T = 2^SH pixels square) from its shared-memory region; compile-time tile size so
// that the pixel loop is unrolled with shifts only; EXTRA = gamma LUT and / or the host-layout copies are wanted
template <int SH, bool EXTRA>
__device__ __forceinline__ void pyr_grad_tile(const FusedArgs& a, int l, const float* __restrict__ R, int S, int hl0, int wl, int hl,
                                              int gx0, int gy0, int tid)
{
	constexpr int T = 1 << SH, N = T * T, ITER = (N + 255) / 256;
	float4* __restrict__ out = a.out[l];
	float* aos = EXTRA ? a.aos[l] : nullptr;
	float* absg = EXTRA ? a.absg[l] : nullptr;
	const float* Blut = EXTRA ? a.Blut : nullptr;
#pragma unroll
	for (int it = 0; it < ITER; it++)
	{
		const int i = tid + 256 * it;
		if (N < 256 && i >= N) break;
		const int py = i >> SH, px = i & (T - 1);
		const int gx = gx0 + px, gy = gy0 + py;
		if (gx >= wl || gy >= hl) continue;
		const float* q = R + (hl0 + py) * S + hl0 + px;
		const float c = q[0];
		const bool interior = gy >= 1 && gy < hl - 1;                 // idx in [wl, wl*(hl-1)), :188
		if (interior && (gx == 0 || gx == wl - 1)) continue;          // written by the row-wrap pass
		float dx, dy, ab;
		grad_px(c, q[-1], q[1], q[-S], q[S], Blut, interior, dx, dy, ab);
		const size_t idx = (size_t)gy * wl + gx;
		out[idx] = make_float4(c, dx, dy, ab);
		if (EXTRA)
		{
			if (aos) { aos[3 * idx] = c; aos[3 * idx + 1] = dx; aos[3 * idx + 2] = dy; }
			if (absg) absg[idx] = ab;
		}
	}
}

template <bool USE_TMA, bool EXTRA>
__global__ void __launch_bounds__(256) pyr_fused_kernel(const __grid_constant__ CUtensorMap tmap, FusedArgs a)
{
	extern __shared__ __align__(128) float sm[];
	__shared__ __align__(8) unsigned long long bar;
	const int tid = threadIdx.x;
	const int x0 = blockIdx.x * TILE, y0 = blockIdx.y * TILE;
	const int S0 = TILE + 2 * a.H;
	if (USE_TMA)
	{
		const unsigned bar_s = (unsigned)__cvta_generic_to_shared(&bar), dst_s = (unsigned)__cvta_generic_to_shared(sm);
		if (tid == 0)
		{
			asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s) : "memory");
			asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
			asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_s), "r"((unsigned)(S0 * S0 * 4)) : "memory");
			asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
			             ::"r"(dst_s), "l"(&tmap), "r"(x0 - a.H), "r"(y0 - a.H), "r"(bar_s) : "memory");
		}
		__syncthreads();      // the barrier is initialised before anybody polls it
		asm volatile(
			"{\n"
			".reg .pred p;\n"
			"PYR_WAIT:\n"
			"mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n"
			"@p bra PYR_DONE;\n"
			"bra PYR_WAIT;\n"
			"PYR_DONE:\n"
			"}\n" ::"r"(bar_s) : "memory");
	}
	else
	{
		for (int i = tid; i < S0 * S0; i += 256)
		{
			const int r = i / S0, c = i - r * S0;
			const int x = x0 - a.H + c, y = y0 - a.H + r;
			sm[i] = (x >= 0 && y >= 0 && x < a.w && y < a.h) ? __ldg(a.src + (size_t)y * a.w + x) : 0.f;
		}
		__syncthreads();
	}
	// ---- level by level: gradients of level l from its shared-memory region (:186-209), then the region of level l+1 by the
	// box cascade (:179-182); the two do not depend on each other, one barrier per level ----
	float* R = sm;
	for (int l = 0; l < a.levels; l++)
	{
		const int S = S0 >> l, T = TILE >> l, hl0 = a.H >> l;
		const int wl = a.w >> l, hl = a.h >> l;
		const int gx0 = x0 >> l, gy0 = y0 >> l;
		float* nxt = R + S * S;
		if (l + 1 < a.levels)
		{
			const int Sd = S >> 1;
			for (int i = tid; i < Sd * Sd; i += 256)
			{
				const int yy = i / Sd, xx = i - yy * Sd;
				const float* p = R + (2 * yy) * S + 2 * xx;
				nxt[i] = fmul(0.25f, fadd(fadd(fadd(p[0], p[1]), p[S]), p[S + 1]));
			}
		}
		switch (l)
		{
		case 0: pyr_grad_tile<6, EXTRA>(a, l, R, S, hl0, wl, hl, gx0, gy0, tid); break;
		case 1: pyr_grad_tile<5, EXTRA>(a, l, R, S, hl0, wl, hl, gx0, gy0, tid); break;
		case 2: pyr_grad_tile<4, EXTRA>(a, l, R, S, hl0, wl, hl, gx0, gy0, tid); break;
		case 3: pyr_grad_tile<3, EXTRA>(a, l, R, S, hl0, wl, hl, gx0, gy0, tid); break;
		case 4: pyr_grad_tile<2, EXTRA>(a, l, R, S, hl0, wl, hl, gx0, gy0, tid); break;
		default: pyr_grad_tile<1, EXTRA>(a, l, R, S, hl0, wl, hl, gx0, gy0, tid); break;
		}
		float4* out = a.out[l];
		float* aos = a.aos[l];
		float* absg = a.absg[l];
		// the reference's dx at x = 0 / w-1 reads the pixel of the neighbouring row (linear index +-1): the CTAs on the
		// image's left / right edge write those pixels with the wrapped neighbour rebuilt from the input
		const bool left = (gx0 == 0), right = (gx0 <= wl - 1 && wl - 1 < gx0 + T);
		for (int side = 0; side < 2; side++)
		{
			if (!(side == 0 ? left : right)) continue;                     // uniform over the CTA
			// the T wrapped neighbours of this tile: column cx of level l, rows ry0 .. ry0 + T - 1.  For l >= 1 they are rebuilt
			// by the whole CTA: the 64 x 2^l strip of the input under that column goes to shared memory (the level-0 region is
			// dead by now) and runs through the same cascade
			const int cx = side == 0 ? wl - 1 : 0, ry0 = side == 0 ? gy0 - 1 : gy0 + 1;
			const float* col = nullptr;
			if (l >= 1)
			{
				__syncthreads();
				float* ping = sm; float* pong = sm + 2048;
				const int Ws = 1 << l;
				for (int i = tid; i < 64 * Ws; i += 256)
				{
					const int r = i >> l, c = i & (Ws - 1);
					const int y = (ry0 << l) + r, x = (cx << l) + c;
					ping[i] = (y >= 0 && y < a.h && x < a.w) ? __ldg(a.src + (size_t)y * a.w + x) : 0.f;
				}
				__syncthreads();
				int wk = Ws, hk = 64;
				for (int k = 0; k < l; k++)
				{
					const int wd = wk >> 1, hd = hk >> 1;
					for (int i = tid; i < wd * hd; i += 256)
					{
						const int yy = i / wd, xx = i - yy * wd;
						const float* p = ping + (2 * yy) * wk + 2 * xx;
						pong[i] = fmul(0.25f, fadd(fadd(fadd(p[0], p[1]), p[wk]), p[wk + 1]));
					}
					__syncthreads();
					float* t = ping; ping = pong; pong = t;
					wk = wd; hk = hd;
				}
				col = ping;                                                // T values: level-l pixel (cx, ry0 + py)
			}
			for (int py = tid; py < T; py += 256)
			{
				const int gx = side == 0 ? 0 : wl - 1, gy = gy0 + py;
				if (gy < 1 || gy >= hl - 1) continue;
				const float wrap = l >= 1 ? col[py] : __ldg(a.src + (size_t)(ry0 + py) * a.w + cx);
				const int px = gx - gx0;
				const float* q = R + (hl0 + py) * S + hl0 + px;
				const float c = q[0];
				// I[idx-1] of x = 0 is the last pixel of the previous row, I[idx+1] of x = w-1 the first of the next (:190)
				const float lf = (side == 0) ? wrap : ((wl > 1) ? q[-1] : wrap);
				const float rt = (side == 1) ? wrap : q[1];
				float dx, dy, ab;
				grad_px(c, lf, rt, q[-S], q[S], a.Blut, true, dx, dy, ab);
				const size_t idx = (size_t)gy * wl + gx;
				out[idx] = make_float4(c, dx, dy, ab);
				if (aos) { aos[3 * idx] = c; aos[3 * idx + 1] = dx; aos[3 * idx + 2] = dy; }
				if (absg) absg[idx] = ab;
			}
		}
		__syncthreads();      // region l+1 complete
		R = nxt;
	}
}

int pyramid_build(dso_b200_ctx* ctx, FrameSlot& fs, const float* color_dev, int w, int h, int levels,
                  const float* d_Blut, bool host_layout)
{
	return pyramid_build_on(ctx, ctx->stream, fs, color_dev, w, h, levels, d_Blut, host_layout);
}

int pyramid_build_on(dso_b200_ctx* ctx, cudaStream_t stream, FrameSlot& fs, const float* color_dev, int w, int h, int levels,
                     const float* d_Blut, bool host_layout)
{
	fs.w = w; fs.h = h; fs.levels = levels;
	static int two_pass = -1;
	if (two_pass < 0) { const char* e = getenv("DSO_B200_K1_TWO_PASS"); two_pass = (e && atoi(e)) ? 1 : 0; }
	if (!two_pass)
	{
		FusedArgs fa;
		memset(&fa, 0, sizeof(fa));
		fa.src = color_dev; fa.w = w; fa.h = h; fa.levels = levels;
		fa.H = levels > 3 ? (1 << (levels - 1)) : 4;
		fa.Blut = d_Blut;
		for (int l = 0; l < levels; l++)
		{
			fa.out[l] = fs.pyr[l];
			fa.aos[l] = host_layout ? ctx->d_aos[l] : nullptr;
			fa.absg[l] = host_layout ? ctx->d_abs[l] : nullptr;
		}
		const int S0 = TILE + 2 * fa.H;
		size_t smem = 0;
		for (int l = 0; l < levels; l++) smem += (size_t)(S0 >> l) * (S0 >> l) * sizeof(float);
		static bool attr_done[64] = { false };
		if (!attr_done[ctx->device & 63])
		{
			CK(cudaFuncSetAttribute(pyr_fused_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
			CK(cudaFuncSetAttribute(pyr_fused_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
			CK(cudaFuncSetAttribute(pyr_fused_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
			CK(cudaFuncSetAttribute(pyr_fused_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
			attr_done[ctx->device & 63] = true;
		}
		// TMA tile load: needs a 16-byte aligned base and pitch; anything else takes the plain-load instantiation
		CUtensorMap tm;
		memset(&tm, 0, sizeof(tm));
		static int no_tma = -1;
		if (no_tma < 0) { const char* e = getenv("DSO_B200_K1_NO_TMA"); no_tma = (e && atoi(e)) ? 1 : 0; }
		bool tma = !no_tma && ((w & 3) == 0) && ((reinterpret_cast<size_t>(color_dev) & 15) == 0);
		if (tma)
		{
			typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
			                              const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
			static encode_fn encode = nullptr;
			static bool looked = false;
			if (!looked)
			{
				void* fn = nullptr;
				cudaDriverEntryPointQueryResult q;
				if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) encode = (encode_fn)fn;
				looked = true;
			}
			if (!encode) tma = false;
			else
			{
				const cuuint64_t gdim[2] = { (cuuint64_t)w, (cuuint64_t)h };
				const cuuint64_t gstr[1] = { (cuuint64_t)w * sizeof(float) };
				const cuuint32_t box[2] = { (cuuint32_t)S0, (cuuint32_t)S0 };
				const cuuint32_t estr[2] = { 1, 1 };
				const CUresult r = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(color_dev), gdim, gstr, box, estr,
				                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
				                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
				if (r != CUDA_SUCCESS) tma = false;
			}
		}
		const dim3 grid((w + TILE - 1) / TILE, (h + TILE - 1) / TILE);
		const bool extra = host_layout || d_Blut != nullptr;
		if (tma) { if (extra) pyr_fused_kernel<true, true><<<grid, 256, smem, stream>>>(tm, fa); else pyr_fused_kernel<true, false><<<grid, 256, smem, stream>>>(tm, fa); }
		else { if (extra) pyr_fused_kernel<false, true><<<grid, 256, smem, stream>>>(tm, fa); else pyr_fused_kernel<false, false><<<grid, 256, smem, stream>>>(tm, fa); }
		LAUNCHED(ctx, 1);
		CK(cudaGetLastError());
		fs.valid = true;
		return 0;
	}
	IntensityArgs ia;
	ia.src = color_dev; ia.w = w; ia.h = h; ia.levels = levels;
	for (int l = 0; l < MAXL; l++) ia.inten[l] = fs.inten[l];
	if (levels > 1)
	{
		dim3 grid((w + TILE - 1) / TILE, (h + TILE - 1) / TILE);
		pyr_intensity_kernel<<<grid, 256, 0, stream>>>(ia);
		LAUNCHED(ctx, 1);
	}
	GradArgs ga;
	memset(&ga, 0, sizeof(ga));
	ga.levels = levels;
	ga.Blut = d_Blut;
	int blocks = 0;
	for (int l = 0; l < levels; l++)
	{
		int wl = w >> l, hl = h >> l;
		GradLevel& L = ga.lv[l];
		L.inten = (l == 0) ? color_dev : fs.inten[l];
		L.out = fs.pyr[l];
		L.aos = host_layout ? ctx->d_aos[l] : nullptr;
		L.absg = host_layout ? ctx->d_abs[l] : nullptr;
		L.w = wl; L.h = hl;
		L.block_start = blocks;
		L.vec = ((wl & 3) == 0) && ((reinterpret_cast<size_t>(L.inten) & 15) == 0);
		int n = wl * hl;
		blocks += (n + 1023) / 1024;   // 256 threads x 4 pixels
	}
	pyr_gradient_kernel<<<blocks, 256, 0, stream>>>(ga);
	LAUNCHED(ctx, 1);
	CK(cudaGetLastError());
	fs.valid = true;
	return 0;
}
