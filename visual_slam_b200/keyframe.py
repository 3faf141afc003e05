"""Key-frame control logic of the tracking loop, same names and results as the reference.

* GetListDiff      src/v2/helper_functions.py:316-326 (called at src/v2/main.py:239) -- on the device
* is_keyframe      the test at src/v2/main.py:221
* cull_due         the test at src/v2/main.py:234 (then Map.DiscardOutlierMapPoints, see mapsoa.MapSoA)
"""
import ctypes as C

import numpy as np
import torch

from . import _cabi


def list_diff_device(kp1, kp2, rowwise=False):
    """kp1 (n1,2), kp2 (n2,2) float64 CUDA tensors -> int32 CUDA tensor of the kept indices of kp1."""
    if kp1.dtype != torch.float64 or kp2.dtype != torch.float64 or not kp1.is_cuda or not kp2.is_cuda:
        raise ValueError("kp1 / kp2 must be float64 CUDA tensors")
    if kp1.dim() != 2 or kp1.shape[1] != 2 or kp2.dim() != 2 or kp2.shape[1] != 2:
        raise ValueError("kp1 / kp2 must have shape (n, 2)")
    kp1, kp2 = kp1.contiguous(), kp2.contiguous()
    n1, n2 = kp1.shape[0], kp2.shape[0]
    keep = torch.empty(max(n1, 1), dtype=torch.uint8, device=kp1.device)
    idx = torch.empty(max(n1, 1), dtype=torch.int32, device=kp1.device)
    n = torch.zeros(1, dtype=torch.int32, device=kp1.device)
    with torch.cuda.device(kp1.device):
        _cabi.check(_cabi.lib().vslam_list_diff(kp1.data_ptr(), n1, kp2.data_ptr(), n2, 1 if rowwise else 0,
                                                keep.data_ptr(), idx.data_ptr(), n.data_ptr(), _cabi.stream_ptr()))
    return idx[:int(n.item())]


def GetListDiff(kp1, kp2, rowwise=False, device="cuda"):
    """Drop-in for the reference's GetListDiff(kp1, kp2): list of the indices of kp1 "not in" kp2.

    rowwise=False reproduces what the reference returns (numpy `x not in kp2`: a key point is dropped when
    any row of kp2 shares either coordinate with it); rowwise=True is the row-wise membership test."""
    a = np.asarray(kp1, dtype=np.float64).reshape(-1, 2)
    b = np.asarray(kp2, dtype=np.float64).reshape(-1, 2)
    out = list_diff_device(torch.from_numpy(a).to(device), torch.from_numpy(b).to(device), rowwise)
    return out.cpu().tolist()


def is_keyframe(i, loop_idx, n_tracked, n_known):
    """main.py:221: more than 20 frames since the last key frame or fewer than 80 tracked map points, and
    fewer than 90 % of the map points the last key frame sees."""
    return (i - loop_idx > 20 or n_tracked < 80) and (n_tracked < 0.9 * n_known)


def cull_due(id_frame):
    """main.py:234: map-point culling runs on every fourth key frame from the sixth on."""
    return id_frame >= 6 and id_frame % 4 == 0
