// C++ consumer of the host mirror (dso-noted-in-chinese_b200/host/dso_b200_host.h): the reference's call sequence
// FrameHessian::makeImages -> CoarseTracker::makeK / setCoarseTrackingRef -> trackNewestCoarse, written as the reference's
// own code would write it, linked against libdso_b200.so.  Scene: a fronto-parallel textured plane at depth 2 seen by a
// camera that moves 0.01 along x, i.e. the second image is the first one shifted by fx * 0.01 / 2 pixels.
// Built and run by tests/test_gpu_host_mirror.py on the GPU box; prints the recovered motion and exits 0 on success.
#include "dso-noted-in-chinese_b200/host/dso_b200_host.h"
#include <vector>

static float texture(float x, float y)
{
	float v = 128.f;
	const float f[6][3] = { {0.071f, 0.013f, 22.f}, {0.017f, 0.083f, 19.f}, {0.044f, -0.052f, 17.f}, {0.11f, 0.031f, 9.f}, {-0.023f, 0.12f, 8.f}, {0.19f, 0.07f, 5.f} };
	for (int i = 0; i < 6; i++) v += f[i][2] * sinf(6.2831853f * (f[i][0] * x + f[i][1] * y) + (float)i);
	return v;
}

int main()
{
	const int w = 640, h = 480, levels = 4;
	const float fx = 525.f, fy = 525.f, cx = 319.5f, cy = 239.5f, depth = 2.f, tx = 0.01f;
	const float shift = fx * tx / depth;                               // pixels
	dso::DeviceContext dev(0, w, h, levels, 4);
	dso::CalibHessian HCalib;
	HCalib.fxl() = fx; HCalib.fyl() = fy; HCalib.cxl() = cx; HCalib.cyl() = cy;
	std::vector<float> img0((size_t)w * h), img1((size_t)w * h);
	for (int y = 0; y < h; y++)
		for (int x = 0; x < w; x++)
		{
			img0[(size_t)y * w + x] = texture((float)x, (float)y);
			img1[(size_t)y * w + x] = texture((float)x + shift, (float)y);   // the camera moved to +x: the scene moves to -x
		}
	dso::FrameHessian ref(&dev, 0), cur(&dev, 1);
	ref.makeImages(img0.data(), 0);
	cur.makeImages(img1.data(), 0);

	// the result of makeCoarseDepthL0 for this scene: every 3rd pixel of every level, constant inverse depth, colour from
	// the reference pyramid the library just returned
	std::vector<std::vector<float> > pu(levels), pv(levels), pid(levels), pc(levels);
	int pc_n[8];
	const float* ptr_u[8]; const float* ptr_v[8]; const float* ptr_id[8]; const float* ptr_c[8];
	for (int l = 0; l < levels; l++)
	{
		const int wl = w >> l, hl = h >> l;
		for (int y = 4; y < hl - 4; y += 3)
			for (int x = 4; x < wl - 4; x += 3)
			{
				pu[l].push_back((float)x); pv[l].push_back((float)y); pid[l].push_back(1.f / depth);
				pc[l].push_back(ref.dIp[l][3 * ((size_t)y * wl + x)]);
			}
		pc_n[l] = (int)pu[l].size();
		ptr_u[l] = pu[l].data(); ptr_v[l] = pv[l].data(); ptr_id[l] = pid[l].data(); ptr_c[l] = pc[l].data();
	}
	dso::CoarseTracker tracker(&dev, 0);
	tracker.makeK(&HCalib);
	tracker.setCoarseTrackingRef(&ref, dso::AffLight(0, 0), pc_n, ptr_u, ptr_v, ptr_id, ptr_c);

	dso::SE3 lastToNew;                                                   // identity prior
	dso::AffLight aff(0, 0);
	dso::Vec5 minRes;
	for (int i = 0; i < 5; i++) minRes[i] = NAN;
	const bool ok = tracker.trackNewestCoarse(&cur, lastToNew, aff, levels - 1, minRes);
	printf("ok=%d t=(%.6f %.6f %.6f) a=%.5f b=%.4f rmse0=%.4f\n", (int)ok, lastToNew.t[0], lastToNew.t[1], lastToNew.t[2], aff.a, aff.b,
	       tracker.lastResiduals[0]);
	// refToNew maps reference-camera points into the new camera: the new camera sits at +tx, so t = (-tx, 0, 0)
	const bool good = ok && fabs(lastToNew.t[0] + tx) < 2e-4 && fabs(lastToNew.t[1]) < 2e-4 && fabs(lastToNew.t[2]) < 5e-4 &&
	                  fabs(lastToNew.R[0] - 1.0) < 1e-5 && fabs(aff.a) < 5e-2 && fabs(aff.b) < 5.0 && tracker.lastResiduals[0] < 1.0;   // gain and offset trade off on a low-contrast texture
	// calcRes / calcGSSSE on the converged state: a positive-definite 8x8 system
	dso::Vec6 r6 = tracker.calcRes(0, lastToNew, aff, 20.f);
	dso::Mat88 H; dso::Vec8 b;
	tracker.calcGSSSE(0, H, b, lastToNew, aff);
	bool diag = true;
	for (int i = 0; i < 8; i++) diag = diag && H(i, i) > 0;
	printf("E=%.3f n=%.0f H00=%.3e\n", r6[0], r6[1], H(0, 0));
	if (!(good && diag && r6[1] > 1000)) { printf("FAILED\n"); return 1; }
	printf("OK\n");
	return 0;
}
