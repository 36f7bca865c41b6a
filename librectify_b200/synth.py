"""Synthetic Manhattan-scene frames (test / benchmark input only — not part of the product path).

Definition follows SURVEY.md §8(d): a square canvas of side 2*max(W,H) with base grey 0.75, 3*L filled
axis-aligned rectangles, L vertical and L horizontal lines, mapped into the frame by a random homography
(canvas corners -> frame corners jittered by +-15 % of W/H), bilinear sampling with reflected border, plus
Gaussian-blurred (sigma 3) texture 0.03*N(0,1) and pixel noise 0.02*N(0,1), clipped to [0, 255/256], float32.

The scene parameters come from numpy ``default_rng(seed)``; warp/noise run in torch on the requested device so the
same function feeds the CPU parity tests (device='cpu', bit-reproducible) and the GPU benchmark (device='cuda').
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F


def _homography(src: np.ndarray, dst: np.ndarray) -> np.ndarray:
    """3x3 H with dst ~ H @ src for four point pairs (rows are (x, y))."""
    A = []
    b = []
    for (x, y), (u, v) in zip(src, dst):
        A.append([x, y, 1, 0, 0, 0, -u * x, -u * y]); b.append(u)
        A.append([0, 0, 0, x, y, 1, -v * x, -v * y]); b.append(v)
    h = np.linalg.solve(np.asarray(A, dtype=np.float64), np.asarray(b, dtype=np.float64))
    return np.append(h, 1.0).reshape(3, 3)


def scene_ops(seed: int, width: int, height: int, n_lines: int = 40, canvas_scale: float = 2.0):
    """The scene as a list of fill operations (y0, y1, x0, x1, grey) on the S x S canvas, plus the homography.
    Returns (S, ops, Hinv 3x3 f64, H 3x3, rng) — the rng is handed on for the noise seed (same draw order as ever)."""
    rng = np.random.default_rng(seed)
    S = int(canvas_scale * max(width, height))
    L = n_lines
    ops = []
    for _ in range(3 * L):
        w = int(rng.integers(S // 40, S // 6))
        h = int(rng.integers(S // 40, S // 6))
        x = int(rng.integers(0, S - w))
        y = int(rng.integers(0, S - h))
        ops.append((y, y + h, x, x + w, np.float32(rng.uniform(0.1, 0.95))))
    for _ in range(L):
        t = int(rng.integers(1, 5)); x = int(rng.integers(0, S - t))
        ops.append((0, S, x, x + t, np.float32(rng.uniform(0.0, 1.0))))
    for _ in range(L):
        t = int(rng.integers(1, 5)); y = int(rng.integers(0, S - t))
        ops.append((y, y + t, 0, S, np.float32(rng.uniform(0.0, 1.0))))
    src = np.array([[0, 0], [S, 0], [S, S], [0, S]], dtype=np.float64)
    dst = np.array([[0, 0], [width, 0], [width, height], [0, height]], dtype=np.float64)
    dst = dst + rng.uniform(-0.15, 0.15, size=(4, 2)) * np.array([width, height])
    H = _homography(src, dst)
    return S, ops, np.linalg.inv(H), H, rng


def scene_canvas(seed: int, width: int, height: int, n_lines: int = 40, canvas_scale: float = 2.0):
    """Draw the axis-aligned canvas and the frame<-canvas homography. Returns (canvas f32 [S,S], Hinv 3x3 f64, H 3x3)."""
    S, ops, Hinv, H, rng = scene_ops(seed, width, height, n_lines, canvas_scale)
    canvas = np.full((S, S), 0.75, dtype=np.float32)
    for y0, y1, x0, x1, v in ops:
        canvas[y0:y1, x0:x1] = v
    return canvas, Hinv, H, rng


def manhattan_frame(seed: int, width: int = 1920, height: int = 1080, n_lines: int = 40,
                    device: str | torch.device = "cpu", canvas_scale: float = 2.0) -> torch.Tensor:
    """One synthetic frame, float32 [height, width] on ``device``."""
    dev = torch.device(device)
    if dev.type == "cuda":      # same canvas, drawn on the device (the fills are exact float32 assignments): no 59 MB upload per frame
        S, ops, Hinv, _, rng = scene_ops(seed, width, height, n_lines, canvas_scale)
        cv = torch.full((S, S), 0.75, dtype=torch.float32, device=dev)
        for y0, y1, x0, x1, v in ops:
            cv[y0:y1, x0:x1] = float(v)
    else:
        canvas, Hinv, _, rng = scene_canvas(seed, width, height, n_lines, canvas_scale)
        S = canvas.shape[0]
        cv = torch.from_numpy(canvas)
    ys, xs = torch.meshgrid(torch.arange(height, device=dev, dtype=torch.float64) + 0.5,
                            torch.arange(width, device=dev, dtype=torch.float64) + 0.5, indexing="ij")
    Hi = torch.from_numpy(Hinv).to(dev)
    X = Hi[0, 0] * xs + Hi[0, 1] * ys + Hi[0, 2]
    Y = Hi[1, 0] * xs + Hi[1, 1] * ys + Hi[1, 2]
    Z = Hi[2, 0] * xs + Hi[2, 1] * ys + Hi[2, 2]
    gx = (X / Z / S * 2 - 1).to(torch.float32)
    gy = (Y / Z / S * 2 - 1).to(torch.float32)
    grid = torch.stack([gx, gy], dim=-1)[None]
    img = F.grid_sample(cv[None, None], grid, mode="bilinear", padding_mode="reflection", align_corners=False)[0, 0]
    g = torch.Generator(device=dev)
    g.manual_seed(int(rng.integers(0, 2**31 - 1)))
    tex = torch.randn((1, 1, height, width), generator=g, device=dev, dtype=torch.float32)
    r = 9
    k = torch.exp(-0.5 * (torch.arange(-r, r + 1, device=dev, dtype=torch.float32) / 3.0) ** 2)
    k = k / k.sum()
    tex = F.conv2d(F.pad(tex, (r, r, 0, 0), mode="reflect"), k.view(1, 1, 1, -1))
    tex = F.conv2d(F.pad(tex, (0, 0, r, r), mode="reflect"), k.view(1, 1, -1, 1))[0, 0]
    tex = tex / tex.std()
    noise = torch.randn((height, width), generator=g, device=dev, dtype=torch.float32)
    img = img + 0.03 * tex + 0.02 * noise
    return img.clamp_(0.0, 255.0 / 256.0).contiguous()


def manhattan_batch(seeds, width=1920, height=1080, n_lines=40, device="cpu") -> torch.Tensor:
    return torch.stack([manhattan_frame(int(s), width, height, n_lines, device) for s in seeds])
