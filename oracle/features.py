"""CPU oracle for feature extraction.  TEST INFRASTRUCTURE ONLY (see
oracle/__init__.py).

Restates reference src/v2/frame.py:5-14 (`FeatureExtractor.compute_features`) in its ORB build variant
(frame.py:8, the variant the Hamming matcher of the hot path consumes):

    pts = cv2.goodFeaturesToTrack(np.mean(img, axis=2).astype(np.uint8), 3000, qualityLevel=0.01, minDistance=7)
    kps = [cv2.KeyPoint(x, y, size=20) ...]            # angle = -1, octave = 0
    kp, des = cv2.ORB_create().compute(img, kps)
    return cv2.KeyPoint_convert(kp), des

The arithmetic lives in OpenCV (third-party, version un-pinned in the reference; 4.13.0 here).  Every step
below restates OpenCV's published algorithm and was pinned bit for bit against cv2 4.13 in the build
container (tests/test_oracle_features.py runs the same comparison live, and tests/golden/features_icl.npz
holds outputs of the reference's own FeatureExtractor on ICL-NUIM frames):

* channel mean, truncated (numpy float64 mean of three uint8 == integer floor of the sum / 3);
* cornerMinEigenVal(blockSize 3, Sobel 3, BORDER_REFLECT_101): Sobel to float32 with the scale
  1/(255*4*3) folded into the smoothing taps and fused multiply-adds in the filter loops (AVX2 build),
  float32 products, 3x3 un-normalised box filter whose row sums are (a+b)+c in float64 and whose column
  sums are a RUNNING float64 sum down each column (sum += new row; out = float32(sum); sum -= old row),
  and (a+c) - sqrt((a-c)^2 + b^2) in float32 without contraction;
* goodFeaturesToTrack: threshold at float32(max * quality), strict 3x3 local maxima (dilate), interior
  pixels only, ordering by (value descending, pixel address descending), greedy minimum-distance
  selection (a corner is kept iff no kept corner of higher order lies at squared distance < d^2), first
  maxCorners kept;
* ORB.compute on provided key points: BGR2GRAY (15-bit fixed point), key points closer than 31 px to the
  border dropped (order kept), 7x7 sigma-2 Gaussian blur of the level image -- ORB blurs a SUB-MATRIX of its
  bordered pyramid buffer, which sends cv2.GaussianBlur past its bit-exact fixed-point branch into the
  float32 separable filter (row: s = k0*p0, then fma(k_i, p_i, s); column: s = k3*c, then
  fma(k_{3+j}, up_j + down_j, s); round-half-even to uint8; BORDER_REFLECT_101) -- rBRIEF with the 256
  learnt test pairs rotated by the key point angle (float32 products, round-half-even), WTA_K = 2,
  LSB-first packing.

The 256 test pairs (`ORB_PATTERN`) are OpenCV's published table (Rublee et al., ICCV 2011); they were
recovered here from cv2's behaviour (tests/golden/make_orb_pattern.py: single-pixel probes through
cv2.ORB.compute with angle 0) and are committed as tests/golden/orb_pattern.npy — the product carries the
same table in visual_slam_b200/orb_pattern.py.
"""
import os

import numpy as np

f32 = np.float32
# cv2.getGaussianKernel(7, 2, CV_32F), as float32 bit patterns (checked against cv2 in the tests)
GAUSS7_BITS = np.array([0x3d8fafb1, 0x3e06387e, 0x3e434a39, 0x3e5d4ae0, 0x3e434a39, 0x3e06387e, 0x3d8fafb1],
                       dtype=np.uint32)
GAUSS7 = GAUSS7_BITS.view(np.float32)
EDGE = 31


def load_pattern():
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "orb_pattern.npy"))


def gray_mean(img):
    """frame.py:11: np.mean(img, axis=2).astype(np.uint8)."""
    return (img.astype(np.int32).sum(axis=2) // 3).astype(np.uint8)


def gray_bgr(img):
    """cv2.cvtColor(BGR2GRAY) on uint8: (B*3735 + G*19235 + R*9798 + 2^14) >> 15."""
    b, g, r = (img[..., i].astype(np.int64) for i in range(3))
    return ((b * 3735 + g * 19235 + r * 9798 + (1 << 14)) >> 15).astype(np.uint8)


def _fma(a, b, c):
    # float32 fused multiply-add: the product of two float32 is exact in float64; the sum is rounded once to
    # float64 and once to float32 (double rounding could differ from a true FMA only on a 2^-29 tie pattern).
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(f32)


def row_vector_width(W):
    """OpenCV's uint8 -> float32 ROW filters (AVX2 build) process 32 pixels per iteration with fused
    multiply-adds; the remaining W % 32 columns run a scalar loop that rounds every product.  (The float32
    column filters fuse in both their vector body and their tail.)  Pinned against cv2 4.13 for odd widths."""
    return (W // 32) * 32


def sobel_pair(gray):
    """cv2.Sobel(gray, CV_32F, 1, 0 / 0, 1, ksize=3, scale=1/3060) as cornerEigenValsVecs calls it."""
    scale = 1.0 / (255.0 * 4.0 * 3.0)
    s1, s2 = f32(1 * scale), f32(2 * scale)
    p = np.pad(gray.astype(f32), 1, mode="reflect")
    # d/dx: row taps (-1, 0, 1) exact, column taps (s1, s2, s1): fma(S0 + S2, s1, fma(S1, s2, 0))
    rx = p[:, 2:] - p[:, :-2]
    dx = _fma(rx[:-2] + rx[2:], s1, (rx[1:-1] * s2).astype(f32))
    # d/dy: row taps (s1, s2, s1) as k0*a, then fma(k1, b, .), fma(k2, c, .); column taps (-1, 0, 1) exact
    a, b, c = p[:, :-2], p[:, 1:-1], p[:, 2:]
    r = _fma(s1, c, _fma(s2, b, (s1 * a).astype(f32)))
    wv = row_vector_width(gray.shape[1])
    if wv < gray.shape[1]:  # scalar tail of the uint8 -> float32 row filter: products rounded, no FMA
        t = ((s1 * a[:, wv:]).astype(f32) + (s2 * b[:, wv:]).astype(f32)).astype(f32)
        r[:, wv:] = (t + (s1 * c[:, wv:]).astype(f32)).astype(f32)
    dy = r[2:] - r[:-2]
    return dx, dy


def _box3_running(a):
    H, W = a.shape
    pp = np.pad(a.astype(np.float64), 1, mode="reflect")
    rs = (pp[:, :-2] + pp[:, 1:-1]) + pp[:, 2:]
    out = np.empty((H, W), f32)
    acc = np.zeros(W)
    acc = acc + rs[0]
    acc = acc + rs[1]
    for y in range(H):
        s0 = acc + rs[y + 2]
        out[y] = s0.astype(f32)
        acc = s0 - rs[y]
    return out


def corner_min_eigen_val(gray):
    """cv2.cornerMinEigenVal(gray, 3, ksize=3), bit for bit."""
    dx, dy = sobel_pair(gray)
    A = _box3_running((dx * dx).astype(f32))
    B = _box3_running((dx * dy).astype(f32))
    C = _box3_running((dy * dy).astype(f32))
    a = (A * f32(0.5)).astype(f32)
    c = (C * f32(0.5)).astype(f32)
    t = (a - c).astype(f32)
    t = ((t * t).astype(f32) + (B * B).astype(f32)).astype(f32)
    return ((a + c).astype(f32) - np.sqrt(t).astype(f32)).astype(f32)


def good_features_to_track(gray, max_corners=3000, quality=0.01, min_distance=7.0, eig=None):
    """cv2.goodFeaturesToTrack(gray, max_corners, quality, min_distance) -> (N, 2) float32 (x, y)."""
    if eig is None:
        eig = corner_min_eigen_val(gray)
    H, W = eig.shape
    thr = f32(float(eig.max()) * quality)
    e = np.where(eig > thr, eig, f32(0))
    pad = np.pad(e, 1, mode="constant", constant_values=-np.inf)
    dil = np.max([pad[dy:dy + H, dx:dx + W] for dy in range(3) for dx in range(3)], axis=0)
    cand = (e != 0) & (e == dil)
    cand[0, :] = cand[-1, :] = False
    cand[:, 0] = cand[:, -1] = False
    ys, xs = np.nonzero(cand)
    vals = e[ys, xs]
    addr = ys * W + xs
    order = np.lexsort((-addr, -vals.astype(np.float64)))
    ys, xs = ys[order], xs[order]
    if min_distance < 1:
        keep = np.arange(len(xs))[:max_corners if max_corners > 0 else None]
        return np.stack([xs[keep], ys[keep]], axis=1).astype(f32)
    cell = int(round(min_distance))
    gw, gh = (W + cell - 1) // cell, (H + cell - 1) // cell
    grid = [[[] for _ in range(gw)] for _ in range(gh)]
    d2 = min_distance * min_distance
    out = []
    for x, y in zip(xs.tolist(), ys.tolist()):
        cx, cy = x // cell, y // cell
        good = True
        for yy in range(max(0, cy - 1), min(gh - 1, cy + 1) + 1):
            for xx in range(max(0, cx - 1), min(gw - 1, cx + 1) + 1):
                for (px, py) in grid[yy][xx]:
                    if (x - px) ** 2 + (y - py) ** 2 < d2:
                        good = False
                        break
                if not good:
                    break
            if not good:
                break
        if good:
            grid[cy][cx].append((x, y))
            out.append((x, y))
            if 0 < max_corners == len(out):
                break
    return np.asarray(out, dtype=f32).reshape(-1, 2)


def gaussian_blur7(gray):
    """The blur ORB applies to its level image (see the module docstring): float32 separable 7x7 with
    fused multiply-adds, == cv2.sepFilter2D(gray, CV_8U, k, k, borderType=BORDER_REFLECT_101)."""
    H, W = gray.shape
    k = GAUSS7
    p = np.pad(gray.astype(f32), 3, mode="reflect")
    h = (k[0] * p[:, 0:W]).astype(f32)
    for i in range(1, 7):
        h = _fma(k[i], p[:, i:i + W], h)
    wv = row_vector_width(W)
    if wv < W:
        t = (k[0] * p[:, wv:W]).astype(f32)
        for i in range(1, 7):
            t = (t + (k[i] * p[:, wv + i:W + i]).astype(f32)).astype(f32)
        h[:, wv:] = t
    v = (k[3] * h[3:3 + H]).astype(f32)
    for j in range(1, 4):
        v = _fma(k[3 + j], (h[3 + j:3 + j + H] + h[3 - j:3 - j + H]).astype(f32), v)
    return np.clip(np.rint(v), 0, 255).astype(np.uint8)


def border_keep(pts, W, H):
    """KeyPointsFilter::runByImageBorder(kps, size, 31): keep 31 <= x < W-31 and 31 <= y < H-31."""
    pts = np.asarray(pts, f32).reshape(-1, 2)
    return (pts[:, 0] >= EDGE) & (pts[:, 0] < W - EDGE) & (pts[:, 1] >= EDGE) & (pts[:, 1] < H - EDGE)


def _round_half_even_f32(v):
    return np.rint(v.astype(f32)).astype(np.int64)


def orb_descriptors(blurred, pts, angles_deg, pattern=None):
    """computeOrbDescriptors at level 0, WTA_K = 2.  pts (N, 2) float32 inside the 31-px border,
    angles_deg (N,) float32 (KeyPoint.angle; -1 for the reference's hand-made key points)."""
    pat = load_pattern() if pattern is None else pattern
    px, py = pat[:, 0].astype(f32), pat[:, 1].astype(f32)
    N = len(pts)
    out = np.zeros((N, 32), np.uint8)
    H, W = blurred.shape
    bp = np.pad(blurred, 32, mode="reflect")
    for n in range(N):
        ang = f32(angles_deg[n]) * f32(np.pi / f32(180.0))
        a, b = f32(np.cos(np.float64(ang))), f32(np.sin(np.float64(ang)))
        cx = int(np.rint(f32(pts[n, 0]))); cy = int(np.rint(f32(pts[n, 1])))
        x = ((px * a).astype(f32) - (py * b).astype(f32)).astype(f32)
        y = ((px * b).astype(f32) + (py * a).astype(f32)).astype(f32)
        ix = _round_half_even_f32(x); iy = _round_half_even_f32(y)
        v = bp[cy + 32 + iy, cx + 32 + ix].astype(np.int32)
        bits = (v[0::2] < v[1::2]).astype(np.uint8)
        out[n] = np.packbits(bits, bitorder="little")
    return out


def compute_features(img, max_corners=3000, quality=0.01, min_distance=7.0, angle=-1.0):
    """frame.py:10-14 with the ORB extractor.  Returns (kp (N, 2) float32, des (N, 32) uint8)."""
    H, W = img.shape[:2]
    pts = good_features_to_track(gray_mean(img), max_corners, quality, min_distance)
    pts = pts[border_keep(pts, W, H)]
    des = orb_descriptors(gaussian_blur7(gray_bgr(img)), pts, np.full(len(pts), angle, f32))
    return pts, des
