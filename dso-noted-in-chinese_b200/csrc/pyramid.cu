// K1 -- image / gradient pyramid, replaces FrameHessian::makeImages
// (reference: FullSystem/HessianBlocks.cpp:137-211).
//
// Two launches per frame, whatever the number of levels:
//   1. pyr_intensity_kernel: each CTA owns a 64x64 tile of the input image and cascades the exact
//      2x2 box sums  0.25f*(((a+b)+c)+d)  (:179-182) through shared memory for all coarser levels,
//      so level l+1 is bit-identical to the reference's level-by-level loop.
//   2. pyr_gradient_kernel: all levels in one grid; each thread handles 4 consecutive pixels with
//      float4 loads of the three rows and writes (I,dx,dy,|grad|^2) as one float4 per pixel (the
//      device-resident layout every gather in the tracker / BA uses) and, on request, the
//      reference's 12-byte AoS + absSquaredGrad arrays through 16-byte stores for the host copy.
// HBM-bound stencil: algorithmic bytes per frame = 4*w*h in + 16*sum_l(w_l*h_l) out.
#include "common.cuh"

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

int pyramid_build(dso_b200_ctx* ctx, FrameSlot& fs, const float* color_dev, int w, int h, int levels,
                  const float* d_Blut, bool host_layout)
{
	return pyramid_build_on(ctx, ctx->stream, fs, color_dev, w, h, levels, d_Blut, host_layout);
}

int pyramid_build_on(dso_b200_ctx* ctx, cudaStream_t stream, FrameSlot& fs, const float* color_dev, int w, int h, int levels,
                     const float* d_Blut, bool host_layout)
{
	fs.w = w; fs.h = h; fs.levels = levels;
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
