"""Seeded synthetic DSO workloads (images, key-frame window, active points, residual graph).

There is no dataset access on the benchmark box, so every input of the hot path is generated here:
a world made of textured planes rendered by exact ray/plane intersection (so ground-truth depth and
pose are known), a key-frame window with ~2000 active points and their host->target residuals
(BASELINE.json configs[1]), and follow-up frames for CoarseTracker::trackNewestCoarse (configs[0]).
Only numpy is used; everything is a pure function of (seed, width, height, counts).
"""
import numpy as np

PATTERN8 = np.array([[0, -2], [-1, -1], [1, -1], [-2, 0], [0, 0], [2, 0], [-1, 1], [0, 2]], dtype=np.int32)
SCALE_XI_ROT, SCALE_XI_TRANS, SCALE_A, SCALE_B = 1.0, 0.5, 10.0, 1000.0
OUTLIER_TH_SUM_COMPONENT = 50.0 * 50.0


# ---------------------------------------------------------------------------------------------
# SE3 helpers (float64).  Pose = (R 3x3, t 3); tangent = (translation[3], rotation[3]).
def hat(w):
    return np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]], dtype=np.float64)


def se3_exp(xi):
    xi = np.asarray(xi, dtype=np.float64)
    u, w = xi[:3], xi[3:]
    th2 = float(w @ w)
    th = np.sqrt(th2)
    if th < 1e-5:
        A = 1.0 - th2 / 6.0
        B = 0.5 - th2 / 24.0
        C = 1.0 / 6.0 - th2 / 120.0
    else:
        A = np.sin(th) / th
        B = (1.0 - np.cos(th)) / th2
        C = (th - np.sin(th)) / (th2 * th)
    W = hat(w)
    W2 = W @ W
    R = np.eye(3) + A * W + B * W2
    V = np.eye(3) + B * W + C * W2
    return R, V @ u


def se3_mul(a, b):
    return a[0] @ b[0], a[0] @ b[1] + a[1]


def se3_inv(a):
    Rt = a[0].T
    return Rt, -(Rt @ a[1])


def pose12(T):
    return np.concatenate([np.asarray(T[0], dtype=np.float64).reshape(9), np.asarray(T[1], dtype=np.float64).reshape(3)])


def from12(p):
    p = np.asarray(p, dtype=np.float64)
    return p[:9].reshape(3, 3).copy(), p[9:12].copy()


def quat_pose(qw, qx, qy, qz):
    """SE3(Quaterniond(w,x,y,z), 0): Sophus normalises the quaternion."""
    q = np.array([qw, qx, qy, qz], dtype=np.float64)
    q /= np.linalg.norm(q)
    w, x, y, z = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    return R, np.zeros(3)


def se3_log(T):
    R, t = T
    c = min(1.0, max(-1.0, (np.trace(R) - 1.0) * 0.5))
    th = np.arccos(c)
    f = 0.5 + th * th / 12.0 if th < 1e-7 else th / (2.0 * np.sin(th))
    w = f * np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    W = hat(w)
    th2 = th * th
    if th < 1e-5:
        k = 1.0 / 12.0
    else:
        A = np.sin(th) / th
        B = (1.0 - np.cos(th)) / th2
        k = (1.0 - A / (2.0 * B)) / th2
    Vinv = np.eye(3) - 0.5 * W + k * (W @ W)
    return np.concatenate([Vinv @ t, w])


# ---------------------------------------------------------------------------------------------
class Scene:
    """Textured planes; camera y points down, z forward.  World -> camera poses."""

    def __init__(self, seed, w=640, h=480, n_gratings=32):
        self.rng = np.random.default_rng(seed)
        self.w, self.h = w, h
        s = w / 640.0
        self.fx = self.fy = 525.0 * s
        self.cx, self.cy = (w - 1) * 0.5, (h - 1) * 0.5
        rng = self.rng
        # wall: roughly fronto-parallel, tilted; floor below the camera
        n1 = np.array([rng.uniform(-0.25, 0.25), rng.uniform(-0.15, 0.15), 1.0])
        n1 /= np.linalg.norm(n1)
        n2 = np.array([rng.uniform(-0.05, 0.05), 1.0, rng.uniform(-0.1, 0.0)])
        n2 /= np.linalg.norm(n2)
        self.planes = [(n1, rng.uniform(2.3, 2.8)), (n2, rng.uniform(0.9, 1.1))]
        self.bases = []
        for n, _ in self.planes:
            a = np.array([1.0, 0, 0]) if abs(n[0]) < 0.9 else np.array([0, 1.0, 0])
            e1 = np.cross(n, a)
            e1 /= np.linalg.norm(e1)
            e2 = np.cross(n, e1)
            self.bases.append((e1, e2))
        # gratings: wavelengths of roughly 20..150 px at depth ~2.5
        wl = rng.uniform(0.1, 0.7, size=n_gratings)
        ang = rng.uniform(0, 2 * np.pi, size=n_gratings)
        self.freq = np.stack([np.cos(ang) / wl, np.sin(ang) / wl], axis=1)
        self.phase = rng.uniform(0, 2 * np.pi, size=(len(self.planes), n_gratings))
        self.amp = rng.uniform(0.3, 1.0, size=n_gratings) * np.sqrt(wl / wl.max())
        self.gain = 2.2 / np.sqrt(np.sum(self.amp ** 2))

    def K(self):
        return self.fx, self.fy, self.cx, self.cy

    def render(self, T_cw, aff=(0.0, 0.0), exposure=1.0):
        """Returns (image float32 [h,w] in irradiance units 0..255, depth float64 [h,w])."""
        R, t = T_cw
        w, h = self.w, self.h
        C = -(R.T @ t)
        xs = (np.arange(w, dtype=np.float64) - self.cx) / self.fx
        ys = (np.arange(h, dtype=np.float64) - self.cy) / self.fy
        X, Y = np.meshgrid(xs, ys)
        dirs_c = np.stack([X, Y, np.ones_like(X)], axis=-1)          # z = 1 -> ray parameter == depth
        dirs_w = dirs_c @ R                                            # R^T applied to each row
        best = np.full((h, w), np.inf)
        tex = np.zeros((h, w))
        for pi, ((n, d), (e1, e2)) in enumerate(zip(self.planes, self.bases)):
            denom = dirs_w @ n
            s = (d - n @ C) / np.where(np.abs(denom) < 1e-12, 1e-12, denom)
            valid = (s > 0.05) & (s < best)
            P = C[None, None, :] + dirs_w * s[..., None]
            a = P @ e1
            b = P @ e2
            val = np.zeros((h, w))
            for k in range(len(self.amp)):
                val += self.amp[k] * np.sin(2 * np.pi * (self.freq[k, 0] * a + self.freq[k, 1] * b) + self.phase[pi, k])
            tex = np.where(valid, val, tex)
            best = np.where(valid, s, best)
        img = 127.5 + 117.5 * np.tanh(self.gain * tex)
        img = exposure * np.exp(aff[0]) * img + aff[1]
        depth = np.where(np.isfinite(best), best, 1e6)
        return img.astype(np.float32), depth

    def trajectory(self, n, step_t=0.02, step_deg=0.5, jitter=0.1, start=None):
        """Constant-velocity world->camera poses with +-jitter on each step."""
        rng = self.rng
        vdir = np.array([1.0, 0.15, 0.35])
        vdir /= np.linalg.norm(vdir)
        wdir = np.array([0.2, 1.0, 0.1])
        wdir /= np.linalg.norm(wdir)
        T = (np.eye(3), np.zeros(3)) if start is None else start
        out = [T]
        for _ in range(n - 1):
            xi = np.concatenate([vdir * step_t, wdir * np.deg2rad(step_deg)])
            xi = xi * (1.0 + rng.uniform(-jitter, jitter, size=6))
            T = se3_mul(se3_exp(xi), T)
            out.append(T)
        return out


def numpy_pyramid(img, levels):
    """Independent float32 box pyramid (used to sanity-check makeImages channel 0)."""
    out = [img.astype(np.float32)]
    for _ in range(1, levels):
        p = out[-1]
        h, w = p.shape[0] // 2, p.shape[1] // 2
        q = np.float32(0.25) * (((p[0:2 * h:2, 0:2 * w:2] + p[0:2 * h:2, 1:2 * w:2]) + p[1:2 * h:2, 0:2 * w:2]) + p[1:2 * h:2, 1:2 * w:2])
        out.append(q.astype(np.float32))
    return out


def pyr_levels_for(w, h, max_levels=6):
    """util/globalCalib.cpp:49-55"""
    lv, wl, hl = 1, w, h
    while wl % 2 == 0 and hl % 2 == 0 and wl * hl > 5000 and lv < max_levels:
        wl //= 2
        hl //= 2
        lv += 1
    return lv


# ---------------------------------------------------------------------------------------------
class Window:
    """Key-frame window in flat arrays (the layout the C ABI takes)."""
    pass


def make_window(seed=0, w=640, h=480, nF=7, n_points=2000, kf_step_t=0.06, kf_step_deg=1.5,
                pose_noise=3e-4, depth_noise=0.005, grad_th=50.0):
    """Builds BASELINE.json configs[1]: nF key frames, ~n_points active points, all host->target residuals."""
    sc = Scene(seed, w, h)
    rng = np.random.default_rng(seed + 1000)
    W = Window()
    W.scene, W.w, W.h, W.nF = sc, w, h, nF
    W.K = sc.K()
    W.poses_true = sc.trajectory(nF, step_t=kf_step_t, step_deg=kf_step_deg)
    W.aff_true = [(float(rng.uniform(-0.05, 0.05)), float(rng.uniform(-3, 3))) for _ in range(nF)]
    W.aff_true[0] = (0.0, 0.0)
    W.images, W.depths = [], []
    for T, aff in zip(W.poses_true, W.aff_true):
        img, dep = sc.render(T, aff)
        W.images.append(img)
        W.depths.append(dep)
    W.ab_exposure = np.ones(nF, dtype=np.float32)
    W.frameEnergyTH = np.full(nF, 8 * 8 * 8, dtype=np.float32)   # HessianBlocks.h:276

    # frame states: evalPT = true pose perturbed; state (unscaled) small, state_zero a little behind state
    W.evalPT = np.zeros((nF, 12))
    W.state = np.zeros((nF, 10))
    W.state_zero = np.zeros((nF, 10))
    for i, T in enumerate(W.poses_true):
        noise = rng.normal(0, pose_noise, size=6) if i > 0 else np.zeros(6)
        Te = se3_mul(se3_exp(noise), T)
        W.evalPT[i] = pose12(Te)
        st = np.zeros(10)
        if i > 0:
            st[:3] = rng.normal(0, 2e-4, 3) / SCALE_XI_TRANS
            st[3:6] = rng.normal(0, 2e-4, 3) / SCALE_XI_ROT
        st[6] = W.aff_true[i][0] / SCALE_A
        st[7] = W.aff_true[i][1] / SCALE_B
        W.state[i] = st
        sz = st.copy()
        if 0 < i < nF - 1:
            sz[:6] -= rng.normal(0, 5e-5, 6)
            sz[6] -= rng.normal(0, 1e-4)
            sz[7] -= rng.normal(0, 1e-5)
        sz[:6] = 0.0   # setStateZero asserts the pose part of state_zero is ~0 (HessianBlocks.cpp:79)
        W.state_zero[i] = sz
    W.calib_delta = rng.normal(0, 1e-4, 4)

    # points: per host frame, strongest-gradient pixel in each cell of a coarse grid
    per_frame = int(np.ceil(n_points / nF))
    us, vs, hosts, idep, idep0, colors, weights = [], [], [], [], [], [], []
    for f in range(nF):
        img = W.images[f]
        gx = np.zeros_like(img)
        gy = np.zeros_like(img)
        gx[:, 1:-1] = 0.5 * (img[:, 2:] - img[:, :-2])
        gy[1:-1, :] = 0.5 * (img[2:, :] - img[:-2, :])
        g2 = gx * gx + gy * gy
        border = 8
        cell = max(2, int(np.sqrt((w - 2 * border) * (h - 2 * border) / (1.5 * per_frame))))   # 1.5x candidates, cut to per_frame below
        cands = []
        for y0 in range(border, h - border - cell + 1, cell):
            for x0 in range(border, w - border - cell + 1, cell):
                blk = g2[y0:y0 + cell, x0:x0 + cell]
                k = int(np.argmax(blk))
                yy, xx = y0 + k // cell, x0 + k % cell
                if g2[yy, xx] > grad_th * (640.0 / w) ** 2:      # the texture's pixel gradients shrink with resolution
                    cands.append((xx, yy))
        rng.shuffle(cands)
        cands = cands[:per_frame]
        for (xx, yy) in cands:
            d_true = 1.0 / W.depths[f][yy, xx]
            d = d_true * (1.0 + rng.normal(0, depth_noise))
            us.append(float(xx))
            vs.append(float(yy))
            hosts.append(f)
            idep.append(d)
            idep0.append(d * (1.0 + rng.normal(0, 0.002)))
            col = np.array([img[yy + dy, xx + dx] for dx, dy in PATTERN8], dtype=np.float32)
            wg = np.array([np.sqrt(OUTLIER_TH_SUM_COMPONENT / (OUTLIER_TH_SUM_COMPONENT + g2[yy + dy, xx + dx]))
                           for dx, dy in PATTERN8], dtype=np.float32)
            colors.append(col)
            weights.append(wg)
    W.nP = len(us)
    W.u = np.array(us, dtype=np.float32)
    W.v = np.array(vs, dtype=np.float32)
    W.host = np.array(hosts, dtype=np.int32)
    W.idepth_scaled = np.array(idep, dtype=np.float32)
    W.idepth_zero_scaled = np.array(idep0, dtype=np.float32)
    W.color = np.ascontiguousarray(np.stack(colors).astype(np.float32))
    W.weights = np.ascontiguousarray(np.stack(weights).astype(np.float32))
    W.priorF = np.zeros(W.nP, dtype=np.float32)
    W.priorF[rng.random(W.nP) < 0.1] = 50.0 * 50.0      # setting_idepthFixPrior on a few "initialiser" points
    W.deltaF = (W.idepth_scaled - W.idepth_zero_scaled).astype(np.float32)

    # residuals: every point to every other frame, in per-point order (EFPoint::residualsAll)
    rp, rt = [], []
    for p in range(W.nP):
        for t in range(nF):
            if t != hosts[p]:
                rp.append(p)
                rt.append(t)
    W.nR = len(rp)
    W.res_point = np.array(rp, dtype=np.int32)
    W.res_target = np.array(rt, dtype=np.int32)
    W.res_state = np.zeros(W.nR, dtype=np.int32)            # ResState::IN
    W.res_energy = np.zeros(W.nR, dtype=np.float64)
    W.res_isLinearized = np.zeros(W.nR, dtype=np.uint8)
    W.res_isActive = np.zeros(W.nR, dtype=np.uint8)
    return W


def tracker_reference_inputs(W, seed=0):
    """Per-point (centerProjectedTo, HdiF) towards the newest key frame, as makeCoarseDepthL0 consumes them."""
    rng = np.random.default_rng(seed + 2000)
    ref = W.nF - 1
    fx, fy, cx, cy = W.K
    Tr = from12(W.evalPT[ref])
    cu, cv, cid, hdi = [], [], [], []
    for p in range(W.nP):
        hst = int(W.host[p])
        if hst == ref:
            continue
        Th = from12(W.evalPT[hst])
        Trel = se3_mul(Tr, se3_inv(Th))
        x = np.array([(W.u[p] - cx) / fx, (W.v[p] - cy) / fy, 1.0])
        d = float(W.idepth_scaled[p])
        pt = Trel[0] @ x + Trel[1] * d
        if pt[2] <= 0:
            continue
        Ku = fx * pt[0] / pt[2] + cx
        Kv = fy * pt[1] / pt[2] + cy
        if not (1.1 < Ku < W.w - 3 and 1.1 < Kv < W.h - 3):
            continue
        cu.append(Ku)
        cv.append(Kv)
        cid.append(d / pt[2])
        hdi.append(1e-3 * rng.uniform(0.5, 2.0))
    return (np.array(cu, dtype=np.float32), np.array(cv, dtype=np.float32),
            np.array(cid, dtype=np.float32), np.array(hdi, dtype=np.float32))


def tracking_sequence(W, n_frames, seed=0, step_t=0.02, step_deg=0.5):
    """Frames following the newest key frame: (image, true refToNew pose, true aff) per frame."""
    rng = np.random.default_rng(seed + 3000)
    sc = W.scene
    ref_true = W.poses_true[-1]
    poses = sc.trajectory(n_frames + 1, step_t=step_t, step_deg=step_deg, start=ref_true)[1:]
    out = []
    for T in poses:
        aff = (float(rng.uniform(-0.05, 0.05)), float(rng.uniform(-3, 3)))
        img, _ = sc.render(T, aff)
        # refToNew with respect to the (perturbed) evaluation pose of the reference key frame
        rel = se3_mul(T, se3_inv(from12(W.evalPT[-1])))
        out.append((img, rel, aff))
    return out


def motion_hypotheses(fh_2_slast, lastF_2_slast):
    """The pose priors FullSystem::trackNewCoarse tries (FullSystem.cpp:311-354): 5 motion models +
    26 rotation perturbations (the rotDelta loop body runs once, rotDelta = 0.02)."""
    inv = se3_inv(fh_2_slast)
    tries = [se3_mul(inv, lastF_2_slast),
             se3_mul(inv, se3_mul(inv, lastF_2_slast)),
             se3_mul(se3_inv(se3_exp(se3_log(fh_2_slast) * 0.5)), lastF_2_slast),
             lastF_2_slast,
             (np.eye(3), np.zeros(3))]
    base = se3_mul(inv, lastF_2_slast)
    r = 0.02
    quats = [(r, 0, 0), (0, r, 0), (0, 0, r), (-r, 0, 0), (0, -r, 0), (0, 0, -r),
             (r, r, 0), (0, r, r), (r, 0, r), (-r, r, 0), (0, -r, r), (-r, 0, r),
             (r, -r, 0), (0, r, -r), (r, 0, -r), (-r, -r, 0), (0, -r, -r), (-r, 0, -r),
             (-r, -r, -r), (-r, -r, r), (-r, r, -r), (-r, r, r), (r, -r, -r), (r, -r, r), (r, r, -r), (r, r, r)]
    for q in quats:
        tries.append(se3_mul(base, quat_pose(1.0, *q)))
    return tries


def immature_scenario(W, host_idx=2, n_points=1500, n_frames=3, seed=0, step_t=0.03, step_deg=0.7):
    """Inputs of FullSystem::traceNewCoarse (FullSystem.cpp:486-530) for one host key frame: candidate pixels (integer,
    textured, away from the border), their true inverse depths, and for each of n_frames new frames the image and
    hostToFrame_KRKi / hostToFrame_Kt / hostToFrame_affine as float32 (true poses, true affine parameters)."""
    rng = np.random.default_rng(seed + 7000)
    sc = W.scene
    fx, fy, cx, cy = W.K
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], dtype=np.float64)
    img = W.images[host_idx]
    gx = np.zeros_like(img); gy = np.zeros_like(img)
    gx[:, 1:-1] = 0.5 * (img[:, 2:] - img[:, :-2]); gy[1:-1, :] = 0.5 * (img[2:, :] - img[:-2, :])
    g2 = gx * gx + gy * gy
    m = 12
    ys, xs = np.nonzero(g2[m:-m, m:-m] > np.percentile(g2, 70))
    sel = rng.choice(len(xs), size=min(n_points, len(xs)), replace=False)
    u = (xs[sel] + m).astype(np.int32); v = (ys[sel] + m).astype(np.int32)
    idepth_true = (1.0 / W.depths[host_idx][v, u]).astype(np.float32)
    T_host = W.poses_true[host_idx]
    a_h, b_h = W.aff_true[host_idx]
    poses = sc.trajectory(n_frames + 1, step_t=step_t, step_deg=step_deg, start=W.poses_true[-1])[1:]
    frames = []
    for T in poses:
        aff = (float(rng.uniform(-0.05, 0.05)), float(rng.uniform(-3, 3)))
        im, _ = sc.render(T, aff)
        R, t = se3_mul(T, se3_inv(T_host))
        KRKi = (K @ R @ np.linalg.inv(K)).astype(np.float32)
        Kt = (K @ t).astype(np.float32)
        # AffLight::fromToVecExposure with unit exposures (NumType.h:176-189)
        a = np.exp(aff[0] - a_h)
        frames.append(dict(image=im, KRKi=np.ascontiguousarray(KRKi), Kt=np.ascontiguousarray(Kt),
                           aff=np.array([a, aff[1] - a * b_h], dtype=np.float32)))
    return dict(host_image=img, u=u, v=v, idepth_true=idepth_true, frames=frames)


def activation_scenario(W, host_idx=2, n_points=800, seed=0, interval=0.05):
    """Inputs of FullSystem::optimizeImmaturePoint (FullSystemOptPoint.cpp:51-209) for the immature points of one host key
    frame: all other key frames of the window as targets with the host's FrameFramePrecalc entries (PRE_RTll, PRE_tTll,
    PRE_aff_mode; true poses and affine parameters, float32) and inverse-depth intervals around the truth; a tenth of the
    points get a wrong interval so that the failure exits are exercised too."""
    rng = np.random.default_rng(seed + 9000)
    sc = immature_scenario(W, host_idx=host_idx, n_points=n_points, n_frames=1, seed=seed)
    T_host = W.poses_true[host_idx]
    a_h, b_h = W.aff_true[host_idx]
    targets = [t for t in range(W.nF) if t != host_idx]
    RT, tT, aff = [], [], []
    for t in targets:
        R, tr = se3_mul(W.poses_true[t], se3_inv(T_host))
        a = np.exp(W.aff_true[t][0] - a_h)
        RT.append(R.astype(np.float32).reshape(-1)); tT.append(tr.astype(np.float32))
        aff.append(np.array([a, W.aff_true[t][1] - a * b_h], dtype=np.float32))
    idt = sc["idepth_true"]
    noise = rng.uniform(-0.3, 0.3, len(idt)).astype(np.float32) * interval
    centre = idt * (1 + noise)
    wrong = rng.random(len(idt)) < 0.1
    centre[wrong] *= rng.uniform(1.5, 3.0, wrong.sum()).astype(np.float32)
    return dict(u=sc["u"], v=sc["v"], host_image=sc["host_image"], targets=targets, images=[W.images[t] for t in targets],
                RTll=np.ascontiguousarray(np.stack(RT)), tTll=np.ascontiguousarray(np.stack(tT)), aff=np.ascontiguousarray(np.stack(aff)),
                idepth_min=(centre * (1 - interval)).astype(np.float32), idepth_max=(centre * (1 + interval)).astype(np.float32),
                idepth_true=idt, wrong=wrong)


def initializer_scenario(W, lvl=1, n_points=1500, seed=0, unit_depth=False):
    """Inputs of CoarseInitializer::calcResAndGS (FullSystem/CoarseInitializer.cpp:370-732) on pyramid level `lvl`: the
    newest key frame plays the first frame, a follow-up frame the new one; points on textured pixels of that level with
    their inverse depths (the truth with a little noise, or 1 as trackFrame sets them before the first snap, :111-126),
    a tenth of them flagged bad, a few with a tiny outlier threshold (energy > outlierTH*20 exit), previous energies, iR."""
    rng = np.random.default_rng(seed + 11000 + lvl)
    img, rel, aff = tracking_sequence(W, 1, seed=seed)[0]
    first = W.images[-1]
    wl, hl = W.w >> lvl, W.h >> lvl
    # level-lvl intensity of the first frame for picking textured pixels (2x2 means, like makeImages)
    I = first.astype(np.float32)
    for _ in range(lvl):
        I = 0.25 * (I[0::2, 0::2] + I[1::2, 0::2] + I[0::2, 1::2] + I[1::2, 1::2])
    gx = np.zeros_like(I); gy = np.zeros_like(I)
    gx[:, 1:-1] = 0.5 * (I[:, 2:] - I[:, :-2]); gy[1:-1, :] = 0.5 * (I[2:, :] - I[:-2, :])
    g2 = gx * gx + gy * gy
    m = 4
    ys, xs = np.nonzero(g2[m:hl - m, m:wl - m] > np.percentile(g2, 60))
    sel = rng.choice(len(xs), size=min(n_points, len(xs)), replace=False)
    u = (xs[sel] + m).astype(np.float32) + np.float32(0.1)          # setFirst stores x + 0.1, y + 0.1 (:779-780)
    v = (ys[sel] + m).astype(np.float32) + np.float32(0.1)
    s = 1 << lvl
    depth = W.depths[-1][np.minimum((v * s).astype(int), W.h - 1), np.minimum((u * s).astype(int), W.w - 1)]
    n = len(u)
    if unit_depth:
        idepth = np.ones(n, dtype=np.float32)
        # the initialiser's scale: unit inverse depth <-> translation scaled by the mean depth
        rel = (rel[0], rel[1] / float(np.mean(depth)))
    else:
        idepth = (1.0 / depth * (1 + rng.normal(0, 0.02, n))).astype(np.float32)
    isGood = (rng.uniform(size=n) > 0.1).astype(np.uint8)
    energy = np.stack([rng.uniform(50, 400, n), rng.uniform(0, 1, n)], axis=1).astype(np.float32)
    outlierTH = np.full(n, 8 * 12 * 12, dtype=np.float32)            # patternNum * setting_outlierTH (:786)
    outlierTH[rng.uniform(size=n) < 0.05] = 0.01
    iR = (idepth * (1 + rng.normal(0, 0.05, n))).astype(np.float32)
    return dict(first=first, new=img, lvl=lvl, u=u, v=v, idepth_new=idepth, isGood=isGood, energy=energy, outlierTH=outlierTH,
                iR=iR, pose12=pose12(rel), aff=(float(aff[0]), float(aff[1])), alphaW=150.0 * 150.0, alphaK=2.5 * 2.5, couplingWeight=1.0)
