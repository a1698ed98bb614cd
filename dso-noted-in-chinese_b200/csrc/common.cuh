// Shared declarations of the dso_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <cmath>
#include <vector>
#include "../../include/dso_b200.h"

#define MAXL DSO_B200_MAX_LEVELS
#define MAXF DSO_B200_MAX_FRAMES

void dso_set_error(const char* fmt, ...);

#define CK(call)                                                                                   \
	do {                                                                                           \
		cudaError_t e_ = (call);                                                                   \
		if (e_ != cudaSuccess) {                                                                   \
			dso_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));    \
			return DSO_B200_E_CUDA;                                                                \
		}                                                                                          \
	} while (0)

#define ARG_CHECK(cond, msg)                                                   \
	do {                                                                       \
		if (!(cond)) { dso_set_error("%s: %s", __func__, msg); return DSO_B200_E_ARG; } \
	} while (0)

// Exactly-rounded float helpers: the masks and counts of the reference are decided by float
// comparisons, so every operation feeding them is a separate IEEE op (never contracted to FMA).
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
// Eigen's unrolled 3-term redux: c0 + (c1 + c2)
__device__ __forceinline__ float dot3e(float a0, float b0, float a1, float b1, float a2, float b2)
{
	return fadd(fmul(a0, b0), fadd(fmul(a1, b1), fmul(a2, b2)));
}

// ---- device pyramid of one frame ------------------------------------------------------------------
// pyr[l]: float4 (I, dx, dy, absSquaredGrad) per pixel -- one 16-byte gather per texel for the tracker
// and BA.  inten[l]: plain float intensity plane (inten[0] is the uploaded image).
struct FrameSlot
{
	int w, h, levels;
	bool valid;
	// cross-stream ordering: every asynchronous build of this slot records `ready` on its stream and bumps `version`;
	// each consumer stream (DSO_B200_STREAM_MAIN / _TRACKER0 / _TRACKER1 / _MAPPING) waits for the event once per version
	cudaEvent_t ready;
	unsigned int version;
	unsigned int seen[4];
	cudaStream_t producer;
	void* base;
	float4* pyr[MAXL];
	float* inten[MAXL];
};

struct TrackerSlot
{
	int levels;
	int w[MAXL], h[MAXL];
	float fx[MAXL], fy[MAXL], cx[MAXL], cy[MAXL];
	float Ki[MAXL][9];
	int pc_n[MAXL];
	float4* pc[MAXL];          // (u, v, idepth, color) per reference point
	void* base;
	size_t cap_points;         // per level capacity
	double ref_a, ref_b;
	float ref_exposure;
	bool valid;
	// last calc_res bookkeeping for calc_gs
	int last_lvl;
	int last_warped_n;          // padded to a multiple of 4 like buf_warped_n
	float* d_sums;              // device: reduced sums of the last evaluation
	float* d_partials;          // device: per-CTA partials
	unsigned int* d_ticket;
	float* h_sums;              // pinned
	// batched tracker workspace
	void* d_track_in;
	void* d_track_out;
	float* d_mc_part;           // multi-cluster tracking: cluster totals [hyp][2][16][S_COUNT]
	unsigned int* d_mc_count;   // and the arrival counter per hypothesis
	void* h_track_in;           // mapped pinned ring of TRACK_IN_RING input blocks, read by the kernel over PCIe (no H2D copy node)
	void* h_track_out;          // mapped pinned: the kernel writes its results here, then h_flags[hyp] = seq
	unsigned int* h_flags;      // mapped pinned, one word per hypothesis
	unsigned int seq;           // sequence number of the last enqueued tracking call with host-visible results (0: none yet)
	unsigned int fetched;       // sequence number of the last call whose results dso_b200_track_fetch handed out
	const void* last_in;        // input block of the last call (dso_b200_track_batched_async with pose NULL re-runs it)
	cudaStream_t stream;        // all work of this tracker slot (CoarseTracker object) runs here
	int consumer;               // DSO_B200_STREAM_TRACKER0 / _TRACKER1
};
// h_track_in / h_track_out / h_flags are rings of TRACK_IN_RING blocks indexed by seq % TRACK_IN_RING, so that up to that
// many tracking calls may be in flight before their results are fetched
#define TRACK_IN_RING 4
#define TRACK_IN_BLOCK 8192

struct BAWorkspace;

struct dso_b200_ctx
{
	int device;
	int sm_count;
	cudaStream_t stream;       // DSO_B200_STREAM_MAIN: makeImages (synchronous and _dev variants)
	cudaStream_t ba_stream;    // DSO_B200_STREAM_MAPPING: windowed BA, immature points (the reference's mapping thread)
	int max_w, max_h;
	int n_slots;
	FrameSlot* slots;
	// staging for host-layout outputs of makeImages
	float* d_aos[MAXL];
	float* d_abs[MAXL];
	float* d_Blut;
	float* d_color;            // landing zone of the uploaded input image
	cudaStream_t upload_stream;   // dso_b200_make_images_async: H2D + pyramid of the next frame while the current one is tracked
	float* d_color_up;            // its own landing zone
	TrackerSlot trk[2];
	BAWorkspace* ba;
	void* coarse_depth;        // CoarseDepthWS (coarse_depth.cu), allocated at the first dso_b200_tracker_set_ref_points
	void* trace;               // TraceWS (trace.cu): the immature points of the window
	void* init_ws;             // InitWS (initializer.cu): grow-only buffers of dso_b200_init_calc_res_and_gs
	dso_b200_settings settings;                       // dso_b200_set_settings
	void* act_scratch; size_t act_scratch_bytes;      // dso_b200_optimize_immature: grow-only device scratch
	int tune_threads, tune_cluster, tune_nclusters;   // dso_b200_track_set_tuning (0 = automatic)
	int last_shape[3];                                // threads, cluster size, clusters per hypothesis of the last tracker launch
	long long launches;                               // incremented with __atomic_fetch_add (two host threads may launch)
	void* d_phase_cycles;      // optional: per-phase cycle counters of the persistent tracker (debug)
};

// kernel-launch accounting: while a thread captures a CUDA graph its launches are only recorded, they count when the
// graph is replayed
extern thread_local bool t_capturing;
extern thread_local long long t_capture_launches;
#define LAUNCHED(ctx, n) do { if (t_capturing) t_capture_launches += (long long)(n); else __atomic_fetch_add(&(ctx)->launches, (long long)(n), __ATOMIC_RELAXED); } while (0)

// sums layout of one tracker evaluation
enum
{
	S_E = 0,          // energy
	S_FLOW_T = 1,     // sumSquaredShiftT
	S_FLOW_RT = 2,    // sumSquaredShiftRT
	S_FLOW_N = 3,     // sumSquaredShiftNum
	S_NTERMS = 4,     // numTermsInE        (exact integer in float: < 2^24)
	S_NSAT = 5,       // numSaturated
	S_NWARP = 6,      // numTermsInWarped (unpadded)
	S_ACC = 8,        // 44 values: upper triangle of the 8x8 (36, row-major r<=c) then J^T r (8)
	S_COUNT = 52
};

int pyramid_build_on(dso_b200_ctx* ctx, cudaStream_t stream, FrameSlot& fs, const float* color_dev, int w, int h, int levels,
                     const float* d_Blut, bool host_layout);
// makes `stream` (consumer id DSO_B200_STREAM_*) wait for the last asynchronous build of the slot, once per build
int frame_slot_wait(dso_b200_ctx* ctx, int frame_slot, cudaStream_t stream, int consumer);
// records that the slot's pyramid has just been enqueued on `stream`
int frame_slot_produced(dso_b200_ctx* ctx, FrameSlot& F, cudaStream_t stream);
void tracker_make_K(TrackerSlot& T, int w, int h, int levels, float fx, float fy, float cx, float cy);   // CoarseTracker::makeK
void coarse_depth_destroy(dso_b200_ctx* ctx);
void trace_destroy(dso_b200_ctx* ctx);
void init_destroy(dso_b200_ctx* ctx);
int pyramid_build(dso_b200_ctx* ctx, FrameSlot& fs, const float* color_dev, int w, int h, int levels,
                  const float* d_Blut, bool host_layout);
int ba_create(dso_b200_ctx* ctx);
void ba_destroy(dso_b200_ctx* ctx);
