// Context management and the K1 entry points of the C ABI (include/dso_b200.h).
#include "common.cuh"
#include <cstdarg>

static thread_local char g_err[512] = "";
thread_local bool t_capturing = false;
thread_local long long t_capture_launches = 0;

void dso_set_error(const char* fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_err, sizeof(g_err), fmt, ap);
	va_end(ap);
}

extern "C" const char* dso_b200_last_error(void) { return g_err; }

static size_t pyr_pixels(int w, int h, int levels)
{
	size_t s = 0;
	for (int l = 0; l < levels; l++) s += (size_t)(w >> l) * (h >> l);
	return s;
}
static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static int create_impl(dso_b200_ctx* ctx, int device, int max_w, int max_h, int n_frame_slots);

extern "C" int dso_b200_create(int device, int max_w, int max_h, int n_frame_slots, dso_b200_ctx** out)
{
	ARG_CHECK(out && max_w >= 16 && max_h >= 16 && n_frame_slots >= 1 && n_frame_slots <= 64, "bad arguments");
	int ndev = 0;
	cudaError_t e = cudaGetDeviceCount(&ndev);
	if (e != cudaSuccess || ndev == 0)
	{
		dso_set_error("no CUDA device available (%s); this library has no CPU fallback", cudaGetErrorString(e));
		return DSO_B200_E_CUDA;
	}
	ARG_CHECK(device >= 0 && device < ndev, "device index out of range");
	CK(cudaSetDevice(device));
	cudaDeviceProp prop;
	CK(cudaGetDeviceProperties(&prop, device));
	if (prop.major < 10)
	{
		dso_set_error("device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
		return DSO_B200_E_CUDA;
	}
	dso_b200_ctx* ctx = new dso_b200_ctx();
	memset(ctx, 0, sizeof(*ctx));
	ctx->device = device;
	ctx->sm_count = prop.multiProcessorCount;
	ctx->max_w = max_w; ctx->max_h = max_h; ctx->n_slots = n_frame_slots;
	dso_b200_settings_default(&ctx->settings);
	int rc = create_impl(ctx, device, max_w, max_h, n_frame_slots);
	if (rc)
	{
		// a half-built context is torn down again (dso_b200_destroy copes with missing pieces); the error text survives
		char keep[512];
		strncpy(keep, g_err, sizeof(keep)); keep[sizeof(keep) - 1] = 0;
		dso_b200_destroy(ctx);
		dso_set_error("%s", keep);
		return rc;
	}
	*out = ctx;
	return 0;
}

static int create_impl(dso_b200_ctx* ctx, int device, int max_w, int max_h, int n_frame_slots)
{
	CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
	CK(cudaStreamCreateWithFlags(&ctx->upload_stream, cudaStreamNonBlocking));
	CK(cudaStreamCreateWithFlags(&ctx->ba_stream, cudaStreamNonBlocking));

	// frame slots: per level a float4 image and (for l >= 1) a float intensity plane
	ctx->slots = new FrameSlot[n_frame_slots];
	memset(ctx->slots, 0, sizeof(FrameSlot) * n_frame_slots);
	for (int s = 0; s < n_frame_slots; s++)
	{
		FrameSlot& F = ctx->slots[s];
		size_t bytes = 0;
		for (int l = 0; l < MAXL; l++)
		{
			size_t n = (size_t)(max_w >> l) * (max_h >> l);
			bytes += align_up(n * sizeof(float4), 256);
			if (l >= 1) bytes += align_up(n * sizeof(float), 256);
		}
		CK(cudaMalloc(&F.base, bytes));
		CK(cudaEventCreateWithFlags(&F.ready, cudaEventDisableTiming));
		char* p = (char*)F.base;
		for (int l = 0; l < MAXL; l++)
		{
			size_t n = (size_t)(max_w >> l) * (max_h >> l);
			F.pyr[l] = (float4*)p; p += align_up(n * sizeof(float4), 256);
			if (l >= 1) { F.inten[l] = (float*)p; p += align_up(n * sizeof(float), 256); }
		}
	}
	// staging: input image, host-layout outputs, LUT
	size_t npx0 = (size_t)max_w * max_h;
	CK(cudaMalloc(&ctx->d_color, align_up(npx0 * sizeof(float), 256)));
	CK(cudaMalloc(&ctx->d_color_up, align_up(npx0 * sizeof(float), 256)));
	for (int l = 0; l < MAXL; l++)
	{
		size_t n = (size_t)(max_w >> l) * (max_h >> l);
		CK(cudaMalloc(&ctx->d_aos[l], align_up(3 * n * sizeof(float), 256)));
		CK(cudaMalloc(&ctx->d_abs[l], align_up(n * sizeof(float), 256)));
	}
	CK(cudaMalloc(&ctx->d_Blut, 256 * sizeof(float)));

	// tracker slots
	for (int t = 0; t < 2; t++)
	{
		TrackerSlot& T = ctx->trk[t];
		T.cap_points = npx0;
		T.consumer = DSO_B200_STREAM_TRACKER0 + t;
		CK(cudaStreamCreateWithFlags(&T.stream, cudaStreamNonBlocking));
		size_t bytes = 0;
		for (int l = 0; l < MAXL; l++) bytes += align_up((size_t)(max_w >> l) * (max_h >> l) * sizeof(float4), 256);
		CK(cudaMalloc(&T.base, bytes));
		char* p = (char*)T.base;
		for (int l = 0; l < MAXL; l++) { T.pc[l] = (float4*)p; p += align_up((size_t)(max_w >> l) * (max_h >> l) * sizeof(float4), 256); }
		CK(cudaMalloc(&T.d_sums, S_COUNT * sizeof(float)));
		CK(cudaMalloc(&T.d_partials, (size_t)ctx->sm_count * 4 * S_COUNT * sizeof(float)));
		CK(cudaMalloc(&T.d_ticket, sizeof(unsigned int)));
		CK(cudaMemset(T.d_ticket, 0, sizeof(unsigned int)));
		CK(cudaHostAlloc((void**)&T.h_sums, S_COUNT * sizeof(float), cudaHostAllocDefault));
		CK(cudaMalloc(&T.d_track_in, 4096 * 2));
		CK(cudaMalloc(&T.d_track_out, sizeof(dso_b200_track_result) * DSO_B200_MAX_HYP));
		CK(cudaMalloc(&T.d_mc_part, (size_t)DSO_B200_MAX_HYP * 2 * 16 * S_COUNT * sizeof(float)));
		CK(cudaMalloc(&T.d_mc_count, DSO_B200_MAX_HYP * sizeof(unsigned int)));
		CK(cudaMemset(T.d_mc_count, 0, DSO_B200_MAX_HYP * sizeof(unsigned int)));
		CK(cudaHostAlloc(&T.h_track_in, TRACK_IN_RING * TRACK_IN_BLOCK, cudaHostAllocMapped));
		CK(cudaHostAlloc(&T.h_track_out, sizeof(dso_b200_track_result) * DSO_B200_MAX_HYP * TRACK_IN_RING, cudaHostAllocMapped));
		CK(cudaHostAlloc((void**)&T.h_flags, sizeof(unsigned int) * DSO_B200_MAX_HYP * TRACK_IN_RING, cudaHostAllocMapped));
		memset(T.h_flags, 0, sizeof(unsigned int) * DSO_B200_MAX_HYP * TRACK_IN_RING);
		T.seq = 0; T.fetched = 0; T.last_in = T.d_track_in;
		CK(cudaMemset(T.d_track_in, 0, 4096 * 2));
		T.last_lvl = -1;
	}
	return ba_create(ctx);
}

extern "C" void dso_b200_destroy(dso_b200_ctx* ctx)
{
	if (!ctx) return;
	cudaSetDevice(ctx->device);
	cudaStream_t all[5] = { ctx->upload_stream, ctx->stream, ctx->ba_stream, ctx->trk[0].stream, ctx->trk[1].stream };
	for (int i = 0; i < 5; i++) if (all[i]) cudaStreamSynchronize(all[i]);
	if (ctx->ba) ba_destroy(ctx);
	coarse_depth_destroy(ctx);
	trace_destroy(ctx);
	init_destroy(ctx);
	if (ctx->slots)
	{
		for (int s = 0; s < ctx->n_slots; s++)
		{
			if (ctx->slots[s].base) cudaFree(ctx->slots[s].base);
			if (ctx->slots[s].ready) cudaEventDestroy(ctx->slots[s].ready);
		}
		delete[] ctx->slots;
	}
	cudaFree(ctx->d_color); cudaFree(ctx->d_color_up);      // cudaFree(NULL) is a no-op
	for (int l = 0; l < MAXL; l++) { cudaFree(ctx->d_aos[l]); cudaFree(ctx->d_abs[l]); }
	cudaFree(ctx->d_Blut);
	if (ctx->d_phase_cycles) cudaFree(ctx->d_phase_cycles);
	cudaFree(ctx->act_scratch);
	for (int t = 0; t < 2; t++)
	{
		TrackerSlot& T = ctx->trk[t];
		cudaFree(T.base); cudaFree(T.d_sums); cudaFree(T.d_partials); cudaFree(T.d_ticket);
		cudaFree(T.d_track_in); cudaFree(T.d_track_out); cudaFree(T.d_mc_part); cudaFree(T.d_mc_count);
		if (T.h_sums) cudaFreeHost(T.h_sums);
		if (T.h_track_in) cudaFreeHost(T.h_track_in);
		if (T.h_track_out) cudaFreeHost(T.h_track_out);
		if (T.h_flags) cudaFreeHost(T.h_flags);
	}
	for (int i = 0; i < 5; i++) if (all[i]) cudaStreamDestroy(all[i]);
	delete ctx;
}

extern "C" void dso_b200_settings_default(dso_b200_settings* s)
{
	s->huberTH = 9; s->outlierTHSumComponent = 50 * 50; s->maxPixSearch = 0.027f; s->minTraceTestRadius = 2;
	s->GNItsOnPointActivation = 3; s->trace_stepsize = 1.0f; s->trace_GNIterations = 3; s->trace_GNThreshold = 0.1f;
	s->trace_extraSlackOnTH = 1.2f; s->trace_slackInterval = 1.5f; s->trace_minImprovementFactor = 2; s->minIdepthH_act = 100;
}
void ba_settings_changed(dso_b200_ctx* ctx);
extern "C" int dso_b200_set_settings(dso_b200_ctx* ctx, const dso_b200_settings* s)
{
	ARG_CHECK(ctx && s, "null argument");
	ARG_CHECK(s->huberTH > 0 && s->outlierTHSumComponent > 0 && s->maxPixSearch > 0, "thresholds must be positive");
	ARG_CHECK(s->minTraceTestRadius >= 0 && s->GNItsOnPointActivation >= 0 && s->GNItsOnPointActivation <= 16 &&
	          s->trace_GNIterations >= 0 && s->trace_GNIterations <= 16, "iteration counts out of range");
	ctx->settings = *s;
	ba_settings_changed(ctx);
	return 0;
}

extern "C" void* dso_b200_stream(dso_b200_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
extern "C" void* dso_b200_stream_for(dso_b200_ctx* ctx, int which)
{
	if (!ctx) return nullptr;
	switch (which)
	{
	case DSO_B200_STREAM_MAIN: return (void*)ctx->stream;
	case DSO_B200_STREAM_TRACKER0: return (void*)ctx->trk[0].stream;
	case DSO_B200_STREAM_TRACKER1: return (void*)ctx->trk[1].stream;
	case DSO_B200_STREAM_MAPPING: return (void*)ctx->ba_stream;
	case DSO_B200_STREAM_UPLOAD: return (void*)ctx->upload_stream;
	}
	return nullptr;
}
extern "C" int dso_b200_sync(dso_b200_ctx* ctx)
{
	ARG_CHECK(ctx, "null context");
	CK(cudaStreamSynchronize(ctx->upload_stream));
	CK(cudaStreamSynchronize(ctx->stream));
	CK(cudaStreamSynchronize(ctx->trk[0].stream));
	CK(cudaStreamSynchronize(ctx->trk[1].stream));
	CK(cudaStreamSynchronize(ctx->ba_stream));
	return 0;
}
extern "C" void* dso_b200_host_alloc(size_t bytes)
{
	void* p = nullptr;
	if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) { dso_set_error("cudaHostAlloc(%zu) failed", bytes); return nullptr; }
	return p;
}
extern "C" void dso_b200_host_free(void* p) { if (p) cudaFreeHost(p); }
extern "C" long long dso_b200_launch_count(dso_b200_ctx* ctx) { return ctx ? __atomic_load_n(&ctx->launches, __ATOMIC_RELAXED) : -1; }

static int check_image_args(dso_b200_ctx* ctx, int frame_slot, const float* color, int w, int h, int levels)
{
	ARG_CHECK(ctx && color, "null argument");
	ARG_CHECK(frame_slot >= 0 && frame_slot < ctx->n_slots, "frame slot out of range");
	ARG_CHECK(w >= 16 && h >= 16 && w <= ctx->max_w && h <= ctx->max_h, "image size out of range");
	ARG_CHECK(levels >= 1 && levels <= MAXL, "levels out of range");
	ARG_CHECK((w >> (levels - 1)) >= 2 && (h >> (levels - 1)) >= 2, "too many levels for this size");
	return 0;
}

static int upload_lut(dso_b200_ctx* ctx, const float* B_lut, const float** d_lut)
{
	*d_lut = nullptr;
	if (B_lut)
	{
		CK(cudaMemcpyAsync(ctx->d_Blut, B_lut, 256 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
		*d_lut = ctx->d_Blut;
	}
	return 0;
}

extern "C" int dso_b200_make_images_dev(dso_b200_ctx* ctx, int frame_slot, const float* color_dev, int w, int h, int levels,
                                        const float* B_lut_host)
{
	int rc = check_image_args(ctx, frame_slot, color_dev, w, h, levels);
	if (rc) return rc;
	CK(cudaSetDevice(ctx->device));
	const float* d_lut;
	rc = upload_lut(ctx, B_lut_host, &d_lut);
	if (rc) return rc;
	rc = pyramid_build(ctx, ctx->slots[frame_slot], color_dev, w, h, levels, d_lut, false);
	if (rc) return rc;
	return frame_slot_produced(ctx, ctx->slots[frame_slot], ctx->stream);
}

int frame_slot_produced(dso_b200_ctx* ctx, FrameSlot& F, cudaStream_t stream)
{
	CK(cudaEventRecord(F.ready, stream));
	F.producer = stream;
	F.version++;
	return 0;
}

int frame_slot_wait(dso_b200_ctx* ctx, int frame_slot, cudaStream_t stream, int consumer)
{
	FrameSlot& F = ctx->slots[frame_slot];
	if (F.seen[consumer] != F.version)
	{
		if (F.producer != stream) CK(cudaStreamWaitEvent(stream, F.ready, 0));
		F.seen[consumer] = F.version;
	}
	return 0;
}

extern "C" int dso_b200_make_images_async(dso_b200_ctx* ctx, int frame_slot, const float* color_host, int w, int h, int levels)
{
	int rc = check_image_args(ctx, frame_slot, color_host, w, h, levels);
	if (rc) return rc;
	CK(cudaSetDevice(ctx->device));
	FrameSlot& F = ctx->slots[frame_slot];
	CK(cudaMemcpyAsync(ctx->d_color_up, color_host, (size_t)w * h * sizeof(float), cudaMemcpyHostToDevice, ctx->upload_stream));
	rc = pyramid_build_on(ctx, ctx->upload_stream, F, ctx->d_color_up, w, h, levels, nullptr, false);
	if (rc) return rc;
	return frame_slot_produced(ctx, F, ctx->upload_stream);
}

extern "C" int dso_b200_make_images(dso_b200_ctx* ctx, int frame_slot, const float* color_host, int w, int h, int levels,
                                    const float* B_lut, float* const* dIp_out, float* const* absgrad_out)
{
	int rc = check_image_args(ctx, frame_slot, color_host, w, h, levels);
	if (rc) return rc;
	CK(cudaSetDevice(ctx->device));
	FrameSlot& F = ctx->slots[frame_slot];
	const float* d_lut;
	rc = upload_lut(ctx, B_lut, &d_lut);
	if (rc) return rc;
	bool want_host = false;
	if (dIp_out) for (int l = 0; l < levels; l++) want_host |= (dIp_out[l] != nullptr);
	if (absgrad_out) for (int l = 0; l < levels; l++) want_host |= (absgrad_out[l] != nullptr);
	CK(cudaMemcpyAsync(ctx->d_color, color_host, (size_t)w * h * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
	rc = pyramid_build(ctx, F, ctx->d_color, w, h, levels, d_lut, want_host);
	if (rc) return rc;
	if (want_host)
	{
		for (int l = 0; l < levels; l++)
		{
			size_t n = (size_t)(w >> l) * (h >> l);
			if (dIp_out && dIp_out[l])
				CK(cudaMemcpyAsync(dIp_out[l], ctx->d_aos[l], 3 * n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
			if (absgrad_out && absgrad_out[l])
				CK(cudaMemcpyAsync(absgrad_out[l], ctx->d_abs[l], n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
		}
	}
	CK(cudaStreamSynchronize(ctx->stream));
	return 0;
}
