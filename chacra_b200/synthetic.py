"""Synthetic trajectories of the BASELINE.json shapes.

A base structure with hydrogens (the 5xnd fixture, 1 629 atoms) is tiled as k rigid copies on a
lattice (one chain id per copy) and animated with smooth low-frequency displacement fields, a small
per-atom jitter and a global rotation per frame.  Generation is a handful of torch tensor ops so that
benchmark inputs can be produced directly in HBM; tests use the same code on the CPU.
"""
from __future__ import annotations

import numpy as np

CHAIN_IDS = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789"


def tile_pdb(src: str, dst: str, copies: int, gap: float = 3.0) -> int:
    """Write `copies` translated copies of the ATOM/HETATM records of `src` (chain ids A, B, ...;
    beyond 62 copies chain ids repeat with a resid offset of 1000 per round).  Returns the atom count."""
    atoms = [ln for ln in open(src) if ln[:6] in ("ATOM  ", "HETATM")]
    xyz = np.array([[float(ln[30:38]), float(ln[38:46]), float(ln[46:54])] for ln in atoms])
    ext = xyz.max(0) - xyz.min(0) + gap
    side = int(np.ceil(copies ** (1.0 / 3.0)))
    out, serial = [], 0
    for c in range(copies):
        ix, iy, iz = c % side, (c // side) % side, c // (side * side)
        shift = np.array([ix, iy, iz]) * ext
        chain = CHAIN_IDS[c % len(CHAIN_IDS)]
        roff = 1000 * (c // len(CHAIN_IDS))
        for ln, p in zip(atoms, xyz):
            serial += 1
            q = p + shift
            resid = int(ln[22:26]) + roff
            out.append("%s%5d%s%s%4d%s%8.3f%8.3f%8.3f%s" % (ln[:6], serial % 100000, ln[11:21], chain, resid, ln[26:30],
                                                             q[0], q[1], q[2], ln[54:].rstrip("\n")))
    with open(dst, "w") as fh:
        fh.write("\n".join(out) + "\nEND\n")
    return len(out)


class FrameGenerator:
    """x_f = R_f (x0 + A * sum_m a_{m,f} d_m sin(k_m.x0 + p_m)) + noise;  float32 [F][N][3]."""

    def __init__(self, xyz0, seed: int, amplitude: float = 0.8, noise: float = 0.03, modes: int = 8, device="cpu",
                 rotate: bool = True, quantize: float = 0.0):
        import torch
        self.torch = torch
        self.device = torch.device(device)
        g = torch.Generator(device="cpu")
        g.manual_seed(int(seed))
        x0 = torch.as_tensor(np.asarray(xyz0, np.float32))
        self.center = x0.mean(0, keepdim=True)
        x0 = x0 - self.center
        wl = 40.0 + 40.0 * torch.rand(modes, generator=g)                    # wavelengths 40-80 A
        kdir = torch.nn.functional.normalize(torch.randn(modes, 3, generator=g), dim=1)
        kvec = kdir * (2 * np.pi / wl)[:, None]
        phase = 2 * np.pi * torch.rand(modes, generator=g)
        ddir = torch.nn.functional.normalize(torch.randn(modes, 3, generator=g), dim=1)
        s = torch.sin(x0 @ kvec.T + phase[None, :])                          # [N][M]
        basis = (s.T[:, :, None] * ddir[:, None, :]).reshape(modes, -1)      # [M][3N]
        self.x0 = x0.to(self.device)
        self.basis = (amplitude * basis).to(self.device)
        self.noise, self.rotate, self.quantize, self.modes = float(noise), rotate, float(quantize), modes
        self.gen = torch.Generator(device=self.device)
        self.gen.manual_seed(int(seed) + 7919)

    def frames(self, n: int):
        torch = self.torch
        a = torch.randn(n, self.modes, generator=self.gen, device=self.device) / np.sqrt(self.modes / 2.0)
        x = (a @ self.basis).reshape(n, -1, 3) + self.x0[None]
        if self.rotate:
            q = torch.nn.functional.normalize(torch.randn(n, 4, generator=self.gen, device=self.device), dim=1)
            w, i, j, k = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
            rot = torch.stack([1 - 2 * (j * j + k * k), 2 * (i * j - k * w), 2 * (i * k + j * w),
                               2 * (i * j + k * w), 1 - 2 * (i * i + k * k), 2 * (j * k - i * w),
                               2 * (i * k - j * w), 2 * (j * k + i * w), 1 - 2 * (i * i + j * j)], 1).reshape(n, 3, 3)
            x = x @ rot.transpose(1, 2)
        x = x + self.center.to(self.device)[None]
        if self.noise:
            x = x + self.noise * torch.randn(x.shape, generator=self.gen, device=self.device)
        if self.quantize:
            x = torch.round(x / self.quantize) * self.quantize
        return x.to(torch.float32).contiguous()


def solvate_pdb(src: str, dst: str, target_atoms: int, spacing: float = 3.1, clash: float = 2.4, pad: float = 6.0,
                seed: int = 5) -> int:
    """Append TIP3 waters (resname TIP3, names OH2 H1 H2, O-H 0.9572 A, H-O-H 104.52 deg) on a jittered lattice around
    the atoms of `src` until the file holds about `target_atoms` atoms.  Waters whose oxygen
    lies within `clash` A of a solute heavy atom are left out.  Waters get chain W and resids 1..9999 recycled with
    segment names W000, W001, ... so that residues stay distinct.  Returns the atom count."""
    from scipy.spatial import cKDTree
    atoms = [ln.rstrip("\n") for ln in open(src) if ln[:6] in ("ATOM  ", "HETATM")]
    xyz = np.array([[float(ln[30:38]), float(ln[38:46]), float(ln[46:54])] for ln in atoms])
    heavy = np.array([not (ln[12:16].strip()[:1] == "H" or (ln[12:16].strip()[:1].isdigit() and ln[12:16].strip()[1:2] == "H"))
                      for ln in atoms])
    n_w = max(0, (target_atoms - len(atoms)) // 3)
    rng = np.random.Generator(np.random.PCG64(seed))
    lo, hi = xyz.min(0) - pad, xyz.max(0) + pad
    # grow the box isotropically until the lattice holds enough sites
    while True:
        n_axis = np.maximum(1, np.floor((hi - lo) / spacing).astype(int))
        if int(np.prod(n_axis)) >= int(1.25 * n_w) + 8:
            break
        lo, hi = lo - 1.0, hi + 1.0
    g = np.stack(np.meshgrid(*[np.arange(k) for k in n_axis], indexing="ij"), -1).reshape(-1, 3)
    sites = lo + (g + 0.5) * spacing + rng.uniform(-0.35, 0.35, (len(g), 3))
    near = cKDTree(xyz[heavy]).query(sites, distance_upper_bound=clash)[0]
    sites = sites[~np.isfinite(near)]
    # keep the sites closest to the solute first (a droplet), so that small targets still wet the surface
    d = cKDTree(xyz[heavy]).query(sites)[0]
    sites = sites[np.argsort(d, kind="stable")[:n_w]]
    half = np.deg2rad(104.52) / 2
    out = list(atoms)
    serial = len(atoms)
    for k, o in enumerate(sites):
        # random orientation: orthonormal frame (u, v), hydrogens at +-half angle around u in the (u, v) plane
        u = rng.normal(size=3); u /= np.linalg.norm(u)
        t = rng.normal(size=3); v = t - u * (t @ u); v /= np.linalg.norm(v)
        h1 = o + 0.9572 * (np.cos(half) * u + np.sin(half) * v)
        h2 = o + 0.9572 * (np.cos(half) * u - np.sin(half) * v)
        resid, seg = k % 9999 + 1, "W%03d" % (k // 9999)
        for name, p, el in (("OH2", o, "O"), ("H1", h1, "H"), ("H2", h2, "H")):
            serial += 1
            out.append("ATOM  %5d %-4s TIP3W%4d    %8.3f%8.3f%8.3f  1.00  0.00      %-4s%2s" % (
                serial % 100000, " " + name if len(name) < 4 else name, resid, p[0], p[1], p[2], seg, el))
    with open(dst, "w") as fh:
        fh.write("\n".join(out) + "\nEND\n")
    return len(out)
