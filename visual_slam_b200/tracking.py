"""Per-frame tracking step, batched over frame pairs: descriptor matching -> correspondence gather ->
RANSAC PnP (reference src/v2/main.py:183-204), entirely on the device."""
import torch

from . import _cabi
from .matching import match_batch
from .pnp import pnp_ransac_batch
from .poses import camera_to_world_batch, guess_batch


def gather_correspondences(keep_q, keep_t, n_keep, map_xyz, kp_cur):
    B, nq, nt = keep_q.shape[0], keep_q.shape[1], kp_cur.shape[1]
    X = torch.empty((B, nq, 3), dtype=torch.float32, device=keep_q.device)
    uv = torch.empty((B, nq, 2), dtype=torch.float32, device=keep_q.device)
    with torch.cuda.device(keep_q.device):
        _cabi.check(_cabi.lib().vslam_gather_correspondences(
            _cabi.ptr(keep_q), _cabi.ptr(keep_t), _cabi.ptr(n_keep), nq, nt, B, _cabi.ptr(map_xyz.contiguous()),
            _cabi.ptr(kp_cur.contiguous()), _cabi.ptr(X), _cabi.ptr(uv), _cabi.stream_ptr()))
    return X, uv


def track_batch(desc_kf, desc_cur, map_xyz, kp_cur, words, K, guess=None, ratio=0.8, cross_check=False,
                thr=8.0, confidence=0.99, prev_c2w=None, want_c2w=False):
    """desc_kf u8 [B,nq,32] (descriptors of the key frame's mapped points), desc_cur u8 [B,nt,32],
    map_xyz f32 [B,nq,3], kp_cur f32 [B,nt,2], words i32 [B,H,4].  The Stage-B guess is either `guess` f64 [B,6]
    (rvec, tvec) or, as the reference builds it (main.py:191-194), derived on the device from `prev_c2w` f64
    [B,4,4], the previous frame's camera->world pose.  Returns the match lists and the PnP result dict (pose f64
    [B,6] world->camera); with want_c2w also `c2w` [B,4,4] and `cam7` [B,7], the camera->world pose of main.py:201-203
    as a matrix and in the (q, t) form the BA arrays hold -- nothing leaves the device."""
    if (guess is None) == (prev_c2w is None):
        raise ValueError("give exactly one of guess and prev_c2w")
    if guess is None:
        guess = guess_batch(prev_c2w)
    m = match_batch(desc_kf, desc_cur, ratio, cross_check)
    X, uv = gather_correspondences(m["keep_q"], m["keep_t"], m["n_keep"], map_xyz, kp_cur)
    out = pnp_ransac_batch(X, uv, m["n_keep"], words, K, guess, thr, confidence)
    out.update(keep_q=m["keep_q"], keep_t=m["keep_t"], n_keep=m["n_keep"], X=X, uv=uv, guess=guess)
    if want_c2w:
        p = camera_to_world_batch(out["pose"])
        out.update(c2w=p["c2w"], cam7=p["cam7"])
    return out
