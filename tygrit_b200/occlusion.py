"""Self-occlusion check on the GPU — mirror of ``TyGrit/checker/occlusion.py``.

``check_self_occlusion`` keeps the reference's signature (link-pose dict in, bool out) so that it drops
into a ``CheckFn`` / ``GazeFn``; ``check_self_occlusion_batch`` is the additive batch form (one verdict
per robot state, e.g. per trajectory waypoint) that the kernel is built for.  Both call
``tg_check_occlusion*`` (``csrc/occlusion.cu``); there is no CPU path.
"""

from __future__ import annotations

import numpy as np
import torch

from . import _native as N
from .engine import _stream_ptr
from .fetch_model import LINK_NAMES


def _device() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("tygrit_b200.occlusion needs a CUDA device (there is no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def _targets(target_center, target_radius, S: int, dev: torch.device) -> torch.Tensor:
    c = torch.as_tensor(np.asarray(target_center, dtype=np.float64) if not isinstance(target_center, torch.Tensor) else target_center)
    c = c.to(device=dev, dtype=torch.float64).reshape(-1, 3)
    if c.shape[0] not in (1, S):
        raise ValueError("target_center must be (3,) or (S, 3)")
    r = torch.as_tensor(np.asarray(target_radius, dtype=np.float64) if not isinstance(target_radius, torch.Tensor) else target_radius)
    r = r.to(device=dev, dtype=torch.float64).reshape(-1, 1).expand(c.shape[0], 1) if r.numel() in (1, c.shape[0]) else None
    if r is None:
        raise ValueError("target_radius must be a scalar or one value per target")
    return torch.cat([c, r], dim=1).contiguous()


def check_self_occlusion(link_poses, target_center, target_radius: float, intrinsics, image_shape) -> bool:
    """Reference signature (``checker/occlusion.py:48-66``).  ``link_poses``: dict link name -> (4,4) f64 in
    base_link; links missing from the dict are skipped, a missing ``head_tilt_link`` gives False."""
    dev = _device()
    frames = np.zeros((1, len(LINK_NAMES), 3, 4), dtype=np.float64)
    present = 0
    for i, name in enumerate(LINK_NAMES):
        T = link_poses.get(name)
        if T is None:
            continue
        frames[0, i] = np.asarray(T, dtype=np.float64)[:3, :4]
        present |= 1 << i
    f = torch.from_numpy(frames).to(dev)
    tg = _targets(target_center, target_radius, 1, dev)
    K = np.ascontiguousarray(np.asarray(intrinsics, dtype=np.float64).reshape(3, 3))
    h, w = int(image_shape[0]), int(image_shape[1])
    out = torch.empty(1, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        N.check(N.lib().tg_check_occlusion_frames(f.data_ptr(), present, tg.data_ptr(), 1, 1, K.ctypes.data, h, w,
                                                  out.data_ptr(), _stream_ptr(dev)), "tg_check_occlusion_frames")
    return bool(out.item())


def check_self_occlusion_batch(q10, target_center, target_radius, intrinsics, image_shape) -> torch.Tensor:
    """(S,) bool CUDA tensor: verdict of ``check_self_occlusion`` for every robot state ``q10[s]`` =
    (torso, 7 arm joints, head pan, head tilt); targets (3,) or (S,3) in base_link, radius scalar or (S,)."""
    dev = q10.device if isinstance(q10, torch.Tensor) and q10.is_cuda else _device()
    q = torch.as_tensor(q10).to(device=dev, dtype=torch.float64).reshape(-1, 10).contiguous()
    S = q.shape[0]
    tg = _targets(target_center, target_radius, S, dev)
    K = np.ascontiguousarray(np.asarray(intrinsics, dtype=np.float64).reshape(3, 3))
    h, w = int(image_shape[0]), int(image_shape[1])
    out = torch.empty(S, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        N.check(N.lib().tg_check_occlusion(q.data_ptr(), tg.data_ptr(), tg.shape[0], S, K.ctypes.data, h, w, out.data_ptr(),
                                           _stream_ptr(dev)), "tg_check_occlusion")
    return out.bool()
