"""Host-side helpers for the z-slab partition (include/mantacuda.h "multi-GPU"): the same split the library uses
(csrc/solver.cu createSolver: rank r owns planes [nz*r/P, nz*(r+1)/P)), so a scene can cut its host arrays into the
per-rank slabs that mf_grid_upload_slab / mf_grid_download_slab take."""
import numpy as np


def zsplit(nz, nranks):
    """plane boundaries, len nranks+1"""
    if nranks < 1 or nz < 2 * nranks and nranks > 1:
        raise ValueError("a grid partitioned over %d GPUs needs at least %d z planes" % (nranks, 2 * nranks))
    return [nz * r // nranks for r in range(nranks + 1)]


def slab_of(array, rank, nranks):
    """view of rank's planes of a (nz, ny, nx[, 3]) array"""
    b = zsplit(array.shape[0], nranks)
    return array[b[rank]:b[rank + 1]]


def assemble(slabs):
    return np.concatenate(list(slabs), axis=0)


def mg_level_sizes(size):
    """vertex counts of the GridMg levels (source/multigrid.cpp:256-276)"""
    sx, sy, sz = size
    out = [(sx, sy, sz)]
    for _ in range(100):
        if sx <= 5 and sy <= 5 and sz <= 5:
            break
        if sx * sy * sz <= 1000:
            break
        sx, sy, sz = (sx + 2) // 2, (sy + 2) // 2, (sz + 2) // 2
        out.append((sx, sy, sz))
    return out


def mg_level_pairs(size, level, rank, nranks):
    """z-pair range of `rank` on a colour-blocked multigrid level (csrc/multigrid.cu lvlPairs); solo levels belong to rank 0"""
    sx, sy, sz = mg_level_sizes(size)[level]
    bz = (sz + 1) // 2
    if nranks == 1:
        return 0, bz
    if sx * sy * sz < (1 << 20) or bz < 2 * nranks:   # multigrid.cu lvlSolo
        return (0, bz) if rank == 0 else (0, 0)
    return bz * rank // nranks, bz * (rank + 1) // nranks
