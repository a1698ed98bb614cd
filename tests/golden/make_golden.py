"""Generates tests/golden/golden_small.npz from the CPU oracle (run from the repo root: python tests/golden/make_golden.py).

The reference ships no fixtures, so these vectors are produced by the oracle restatement, which
tests/test_oracle_vs_reference.py pins bit for bit against the reference's own sources (oracle/refbuild); they freeze
its outputs so that neither the oracle nor the CUDA path can drift unnoticed, also where /root/reference is absent.  Inputs are regenerated from the seed; a checksum of the rendered images guards against generator drift.
"""
import os
import sys
import zlib
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g  # noqa: E402
from helpers import TrackScenario  # noqa: E402

SEED, WIDTH, HEIGHT, NF, NPTS = 5, 320, 240, 4, 400


def build(pkg, O):
    W = pkg.synthetic.make_window(seed=SEED, w=WIDTH, h=HEIGHT, nF=NF, n_points=NPTS)
    scen = TrackScenario(pkg, O, W, n_frames=2)
    return W, scen


def compute(pkg, O):
    W, scen = build(pkg, O)
    out = {"image_crc": np.array([zlib.crc32(np.ascontiguousarray(im).tobytes()) for im in W.images], dtype=np.uint32)}
    d, a = O.make_images(W.images[0], scen.levels)
    for l in range(scen.levels):
        out["pyr_crc_%d" % l] = np.array([zlib.crc32(d[l].tobytes()), zlib.crc32(a[l].tobytes())], dtype=np.uint32)
    out["pc_n"] = np.array([len(c[0]) for c in scen.clouds], dtype=np.int32)
    scen.oracle_set_frame(0)
    pose = scen.poses(0)[2]
    aa, bb = scen.frames[0][2]
    out["eval_pose"] = pose
    out["eval_aff"] = np.array([aa, bb])
    for lvl in range(scen.levels):
        r = scen.trk.calc_res(lvl, pose, aa, bb, 20.0)
        H, b = scen.trk.calc_gs(lvl, aa, bb)
        st = scen.trk.get_point_status(len(scen.clouds[lvl][0]))
        out["res_%d" % lvl] = r
        out["H_%d" % lvl] = H
        out["b_%d" % lvl] = b
        out["status_crc_%d" % lvl] = np.array([zlib.crc32(st.tobytes())], dtype=np.uint32)
    ident = pkg.synthetic.pose12((np.eye(3), np.zeros(3)))
    tr = scen.trk.track(ident, scen.a0, scen.b0, scen.levels - 1)
    out["track_pose"] = np.array(tr.pose)
    out["track_aff"] = np.array([tr.aff_a, tr.aff_b])
    out["track_trace"] = np.array(list(tr.trace[:tr.trace_len]), dtype=np.uint8)
    out["track_res"] = np.array(list(tr.lastResiduals))
    pyr = [O.make_images(img, 1)[0][0] for img in W.images]
    B = O.BA(W, pyr)
    lin = B.linearize()
    act, jp = B.apply_res()
    A = B.accumulate_A(False)
    S = B.accumulate_SC(False)
    out["ba_newState"] = lin["newState"].astype(np.int8)
    out["ba_energy"] = lin["newEnergy"]
    out["ba_total"] = np.array([lin["total"]])
    out["ba_J_crc"] = np.array([zlib.crc32(lin["J"][lin["newState"] != 1].tobytes())], dtype=np.uint32)
    out["ba_HA"], out["ba_bA"], out["ba_HS"], out["ba_bS"] = A["H"], A["b"], S["H"], S["b"]
    out["ba_HdiF"] = S["HdiF"]
    return out


if __name__ == "__main__":
    pkg = g.load_package()
    O = g.load_oracle()
    out = compute(pkg, O)
    path = os.path.join(ROOT, "tests", "golden", "golden_small.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")
