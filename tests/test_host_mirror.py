"""The C++ mirror of the reference's class surface compiles against the C ABI header (CPU check)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_host_mirror_compiles(tmp_path):
    src = tmp_path / "m.cpp"
    src.write_text('#include "dso-noted-in-chinese_b200/host/dso_b200_host.h"\n'
                   'int use(dso::CoarseTracker* t, dso::FrameHessian* f) { dso::SE3 T; dso::AffLight a; dso::Vec5 m; '
                   'for (int i = 0; i < 5; i++) m[i] = NAN; return t->trackNewestCoarse(f, T, a, 3, m) ? 1 : 0; }\n'
                   'double ba(dso::EnergyFunctional* ef, const dso::WindowArrays& W, dso::CalibHessian* c) { ef->setWindow(W, c); '
                   'double e = ef->linearizeAll(true); dso::MatXX H; dso::VecX b; ef->accumulateAF_MT(H, b, true); ef->accumulateLF_MT(H, b, true); '
                   'ef->accumulateSCF_MT(H, b, true); double z[4] = {0, 0, 0, 0}; ef->backupStatePoints(); ef->solveSystemF(0, 1e-5, 128 | 2048, z); '
                   'float s0, s1; ef->doStepFromBackupPoints(1.f, s0, s1); return e + ef->lastX[0] + H(0, 0) + b[0]; }\n')
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Werror", "-c", "-I", ROOT, str(src), "-o", str(tmp_path / "m.o")])
