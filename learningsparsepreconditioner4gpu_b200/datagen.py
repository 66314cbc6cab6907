"""Self-contained synthetic system generators for the five BASELINE.json configurations.

The reference's generators (datagen/heat.py, poisson3d_tetmesh.py, elast_twist.py) need tetgen,
pymathprim.geometry and pyssim, none of which exist here, so the same *shape
classes* are produced by a P1 finite-element assembly on a (jittered) Kuhn triangulation of a
structured box: 6 tets per cube, 15-point stencil (self, 6 axis, 6 face-diagonal, 2 body-diagonal
neighbours) -- the same ~15 stored entries per row the reference's TetWild meshes have.  Everything is
written with torch ops so that it runs on the CPU for the parity tests and on the GPU for the 10 M-DOF
bench system (the generator is outside every timed region).

Output contract (what neural_cg/data.py:442-450,500-508 hands to the model): ``edge_index`` int64
``[2,E]`` row-major with ascending columns, diagonal included, structural zeros KEPT in the edge list;
``values`` fp64 ``[E]`` (scalar) or ``[E,3,3]`` (elasticity blocks).

``generate_spd_sparse_matrix`` follows datagen/synthetic.py:10-27 (A = (DM)^T(DM) + alpha*I).
"""
from __future__ import annotations

import itertools
import math
from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np
import torch

# 15 stencil offsets of the Kuhn triangulation, (dx, dy, dz)
_POS = [(1, 0, 0), (0, 1, 0), (0, 0, 1), (1, 1, 0), (0, 1, 1), (1, 0, 1), (1, 1, 1)]
_OFFSETS = [(0, 0, 0)] + _POS + [(-a, -b, -c) for (a, b, c) in _POS]


@dataclass
class GridSystem:
    """One assembled system on an nx*ny*nz node grid (z fastest)."""

    shape: Tuple[int, int, int]
    pos: torch.Tensor          # [N,3] fp64 node coordinates
    edge_index: torch.Tensor   # [2,E] int64
    values: torch.Tensor       # [E] or [E,b,b] fp64
    block_size: int
    node_field: Optional[torch.Tensor] = None  # [N,k] extra node features (coefficient field, deformation)

    @property
    def num_nodes(self) -> int:
        return self.pos.shape[0]

    @property
    def xyz(self) -> torch.Tensor:
        """Coordinates scaled to the unit box (the node feature the reference feeds, heat.py:75-76)."""
        return self.pos / max(self.shape[0] - 1, 1)


def _sorted_offsets(ny: int, nz: int):
    """Stencil offsets sorted by linear-index delta, so that slots are column-ascending."""
    key = lambda o: (o[0] * ny + o[1]) * nz + o[2]
    offs = sorted(_OFFSETS, key=key)
    return offs, {o: s for s, o in enumerate(offs)}


def grid_positions(nx, ny, nz, jitter=0.0, seed=0, device="cpu", extent=None) -> torch.Tensor:
    """Node coordinates; interior nodes are displaced by ``jitter``*h (uniform) so that no stencil
    entry cancels to an exact zero."""
    # default spacing h = 1 exactly: un-jittered grids then assemble to exact zeros on the 8 diagonal
    # stencil slots (the matrix is rescaled by 1/mean|A| downstream, so the length unit is immaterial)
    ext = extent or (float(nx - 1), float(ny - 1), float(nz - 1))
    ax = [torch.linspace(0.0, ext[d], n, dtype=torch.float64, device=device) for d, n in enumerate((nx, ny, nz))]
    X, Y, Z = torch.meshgrid(*ax, indexing="ij")
    pos = torch.stack([X, Y, Z], dim=-1).reshape(-1, 3)
    if jitter > 0:
        g = torch.Generator(device="cpu").manual_seed(seed)
        h = ext[0] / max(nx - 1, 1)
        # generated in chunks on the host RNG for reproducibility across devices
        noise = (torch.rand(pos.shape, generator=g, dtype=torch.float64) - 0.5).to(device) * (2 * jitter * h)
        interior = torch.ones(nx, ny, nz, dtype=torch.bool, device=device)
        for d, n in enumerate((nx, ny, nz)):
            idx = [slice(None)] * 3
            idx[d] = 0
            interior[tuple(idx)] = False
            idx[d] = n - 1
            interior[tuple(idx)] = False
        pos = pos + noise * interior.reshape(-1, 1).to(torch.float64)
    return pos


def _cube_corner_ids(nx, ny, nz, corner, device):
    """Linear node id of ``corner`` (cx,cy,cz in {0,1}) of every cube, flattened in cube order."""
    ix = torch.arange(nx - 1, device=device) + corner[0]
    iy = torch.arange(ny - 1, device=device) + corner[1]
    iz = torch.arange(nz - 1, device=device) + corner[2]
    return ((ix.view(-1, 1, 1) * ny + iy.view(1, -1, 1)) * nz + iz.view(1, 1, -1)).reshape(-1)


def _tet_gradients(p0, p1, p2, p3):
    """Barycentric gradients [T,4,3] and volumes [T] of tets given vertex coordinates [T,3]."""
    a, b, c = p1 - p0, p2 - p0, p3 - p0
    # rows of inv([a b c]^T) by cofactors: g1 = (b x c)/det, g2 = (c x a)/det, g3 = (a x b)/det
    bc = torch.linalg.cross(b, c)
    ca = torch.linalg.cross(c, a)
    ab = torch.linalg.cross(a, b)
    det = (a * bc).sum(-1)
    g1, g2, g3 = bc / det[:, None], ca / det[:, None], ab / det[:, None]
    g0 = -(g1 + g2 + g3)
    return torch.stack([g0, g1, g2, g3], dim=1), det.abs() / 6.0


def assemble_kuhn(
    nx, ny, nz, pos, kind="poisson", kappa_node=None, mass_coeff=0.0, lame=None
) -> Tuple[torch.Tensor, torch.Tensor]:
    """P1 assembly into a dense stencil array.

    Returns ``S`` of shape ``[N,15]`` (poisson) or ``[N,15,3,3]`` (elasticity), slots ordered by
    ascending column, and the lumped mass ``[N]``.
    """
    dev = pos.device
    N = nx * ny * nz
    offs, slot = _sorted_offsets(ny, nz)
    b = 1 if kind == "poisson" else 3
    S = torch.zeros((N, 15) if b == 1 else (N, 15, 3, 3), dtype=torch.float64, device=dev)
    lumped = torch.zeros(N, dtype=torch.float64, device=dev)
    Sf = S.view(N * 15, -1)
    eye3 = torch.eye(3, dtype=torch.float64, device=dev)
    for perm in itertools.permutations(range(3)):
        corners = [(0, 0, 0)]
        cur = [0, 0, 0]
        for axis in perm:
            cur[axis] = 1
            corners.append(tuple(cur))
        ids = [_cube_corner_ids(nx, ny, nz, c, dev) for c in corners]
        grads, vol = _tet_gradients(*[pos[i] for i in ids])
        w = vol
        if kappa_node is not None:
            w = vol * (kappa_node[ids[0]] + kappa_node[ids[1]] + kappa_node[ids[2]] + kappa_node[ids[3]]) / 4.0
        for i in range(4):
            lumped.index_add_(0, ids[i], vol / 4.0)
            for j in range(4):
                d = tuple(corners[j][k] - corners[i][k] for k in range(3))
                flat = ids[i] * 15 + slot[d]  # unique per cube for fixed (perm,i,j)
                gi, gj = grads[:, i], grads[:, j]
                if b == 1:
                    Sf[flat, 0] += w * (gi * gj).sum(-1)
                else:
                    lam, mu = lame
                    blk = lam * gi[:, :, None] * gj[:, None, :] + mu * gi[:, None, :] * gj[:, :, None]
                    blk = blk + mu * (gi * gj).sum(-1)[:, None, None] * eye3
                    Sf[flat] += (w[:, None, None] * blk).reshape(-1, 9)
    if mass_coeff != 0.0:
        s0 = slot[(0, 0, 0)]
        if b == 1:
            S[:, s0] += mass_coeff * lumped
        else:
            S[:, s0] += (mass_coeff * lumped)[:, None, None] * eye3
    return S, lumped


def stencil_to_edges(nx, ny, nz, S) -> Tuple[torch.Tensor, torch.Tensor]:
    """Drop the out-of-grid slots; keep (possibly exactly zero) in-grid entries."""
    dev = S.device
    offs, _ = _sorted_offsets(ny, nz)
    N = nx * ny * nz
    ids = torch.arange(N, device=dev)
    iz = ids % nz
    iy = (ids // nz) % ny
    ix = ids // (ny * nz)
    valid = torch.empty((N, 15), dtype=torch.bool, device=dev)
    cols = torch.empty((N, 15), dtype=torch.int64, device=dev)
    for s, (dx, dy, dz) in enumerate(offs):
        valid[:, s] = (ix + dx >= 0) & (ix + dx < nx) & (iy + dy >= 0) & (iy + dy < ny) & (iz + dz >= 0) & (iz + dz < nz)
        cols[:, s] = ids + (dx * ny + dy) * nz + dz
    rows = ids.view(-1, 1).expand(N, 15)
    edge_index = torch.stack([rows[valid], cols[valid]], dim=0)
    return edge_index, S[valid]


def smooth_field(pos, seed=0, modes=4):
    """Smooth random field in [0,1] (stand-in for the gstools field of datagen/heat.py:82-89)."""
    rng = np.random.RandomState(seed)
    f = torch.zeros(pos.shape[0], dtype=torch.float64, device=pos.device)
    pos = pos / pos.max().clamp(min=1.0)
    for _ in range(modes):
        k = torch.tensor(rng.uniform(1.0, 6.0, 3), dtype=torch.float64, device=pos.device)
        ph = float(rng.uniform(0, 2 * math.pi))
        f += torch.sin(pos @ k + ph)
    f = f - f.min()
    return f / f.max().clamp(min=1e-30)


def poisson3d(nx, ny=None, nz=None, jitter=0.2, seed=0, device="cpu") -> GridSystem:
    """Config C2 / C5: P1 Laplacian on a jittered Kuhn tet grid (poisson3d_tetmesh stand-in)."""
    ny, nz = ny or nx, nz or nx
    pos = grid_positions(nx, ny, nz, jitter, seed, device)
    S, _ = assemble_kuhn(nx, ny, nz, pos)
    ei, v = stencil_to_edges(nx, ny, nz, S)
    return GridSystem((nx, ny, nz), pos, ei, v, 1)


def heat3d(nx, ny=None, nz=None, jitter=0.2, seed=42, var=0.99, mass=1e-4, device="cpu") -> GridSystem:
    """Config C1: L(kappa) + 1e-4*M with kappa = field*var + (1-var) (datagen/heat.py:82-89 shape)."""
    ny, nz = ny or nx, nz or nx
    pos = grid_positions(nx, ny, nz, jitter, seed, device)
    field = smooth_field(pos, seed)
    kappa = field * var + (1.0 - var)
    S, _ = assemble_kuhn(nx, ny, nz, pos, kappa_node=kappa, mass_coeff=mass)
    ei, v = stencil_to_edges(nx, ny, nz, S)
    return GridSystem((nx, ny, nz), pos, ei, v, 1, node_field=field.view(-1, 1))


def elasticity3d(nx=64, ny=16, nz=16, jitter=0.0, seed=0, E=3e6, nu=0.4, rho=1.0, dt=0.01, device="cpu") -> GridSystem:
    """Config C3: linear-elasticity stiffness + rho/dt^2 lumped mass, 3x3 blocks (elast_twist stand-in)."""
    pos = grid_positions(nx, ny, nz, jitter, seed, device)
    lam = E * nu / ((1 + nu) * (1 - 2 * nu))
    mu = E / (2 * (1 + nu))
    S, _ = assemble_kuhn(nx, ny, nz, pos, kind="elasticity", mass_coeff=rho / (dt * dt), lame=(lam, mu))
    ei, v = stencil_to_edges(nx, ny, nz, S)
    g = torch.Generator().manual_seed(seed)
    deform = (torch.randn(pos.shape, generator=g, dtype=torch.float64) * 1e-2).to(device)
    return GridSystem((nx, ny, nz), pos, ei, v, 3, node_field=deform)


def generate_spd_sparse_matrix(n, sparsity=3e-4, condition_amplifier=1e-5, random_state=None):
    """Config C4: A = (D M)^T (D M) + alpha I, M uniform-random sparse in [-1,1], D = diag(linspace(1,alpha,n)).

    Follows datagen/synthetic.py:10-27 (same scipy calls in the same order, so the same
    ``random_state`` reproduces the same matrix).
    """
    import scipy.sparse as sp

    rng = np.random.default_rng(random_state)
    M = sp.random(n, n, density=sparsity, format="csr", random_state=rng)
    M.data = 2.0 * (M.data - 0.5)
    M = sp.diags(np.linspace(1, condition_amplifier, n)) @ M
    A = M.T @ M
    A += condition_amplifier * sp.eye(n)
    return A


def scipy_to_edges(A) -> Tuple[torch.Tensor, torch.Tensor]:
    """COO of a scipy matrix in row-major/col-sorted order (what FolderDataset.load produces for .mtx)."""
    from scipy.sparse import csr_matrix

    csr = csr_matrix(A).sorted_indices()
    coo = csr.tocoo()
    ei = torch.from_numpy(np.vstack([coo.row, coo.col]).astype(np.int64))
    return ei, torch.from_numpy(coo.data.astype(np.float64))


def dirichlet_patch_mask(num_nodes, edge_index, boundary_ids, seed=42, rings=2, seed_count=1):
    """mask[N] = 0 on a patch grown ``rings`` graph rings around ``seed_count`` random boundary nodes
    (datagen_helper.py:140-178 + poisson3d_tetmesh.yaml:16-17), 1 elsewhere."""
    rng = np.random.RandomState(seed)
    pick = rng.choice(len(boundary_ids), size=seed_count, replace=False)
    dev = edge_index.device
    hit = torch.zeros(num_nodes, dtype=torch.bool, device=dev)
    hit[torch.as_tensor(np.asarray(boundary_ids)[pick], device=dev)] = True
    for _ in range(rings):
        nxt = hit.clone()
        nxt[edge_index[1][hit[edge_index[0]]]] = True
        nxt[edge_index[0][hit[edge_index[1]]]] = True
        hit = nxt
    return (~hit).to(torch.float64)


def boundary_node_ids(nx, ny, nz) -> np.ndarray:
    ids = np.arange(nx * ny * nz).reshape(nx, ny, nz)
    m = np.zeros((nx, ny, nz), dtype=bool)
    m[0], m[-1], m[:, 0], m[:, -1], m[:, :, 0], m[:, :, -1] = True, True, True, True, True, True
    return ids[m]


def clamped_ends_mask(nx, ny, nz) -> torch.Tensor:
    """[N,3] mask with both x end faces clamped (datagen/elast_twist.py:59-66)."""
    m = torch.ones(nx, ny, nz, 3, dtype=torch.float64)
    m[0] = 0.0
    m[-1] = 0.0
    return m.reshape(-1, 3)
