"""Restart files: host mirror of the reference's dump / restart path over the C ABI.

Reference: the restart dump in the main loop (main/src/sphexa/sphexa.cpp:232-246), the attribute set of
ParticlesData::loadOrStoreAttributes (sph/include/sph/particles_data.hpp:142-193), Box::loadOrStore
(domain/include/cstone/sfc/box.hpp:168-175), TurbulenceData::loadOrStore (hydro_turb/turbulence_data.hpp:90-125)
and the restart reader (main/src/init/file_init.hpp:93-140): every rank reads an equal contiguous slice of each
field and the first domain sync redistributes the particles.

The container is the plain binary format of csrc/checkpoint.cu (HDF5 is not available), one file per rank:
`<path>` for a single rank, `<path>.<rank>` otherwise.  A run restarted on the same number of ranks continues
bit-identically; on a different number it continues from the same particle set with another summation partition.
"""
import ctypes as C
import os
import numpy as np

from . import binding as B
from .binding import lib, check, vp


def rank_path(path, rank, num_ranks):
    return path if num_ranks == 1 else f"{path}.{rank}"


def find_files(path):
    """files of one dump in rank order: `path` itself or path.0, path.1, ..."""
    if os.path.exists(path):
        return [path]
    out = []
    while os.path.exists(f"{path}.{len(out)}"):
        out.append(f"{path}.{len(out)}")
    if not out:
        raise FileNotFoundError(path)
    # the dump says how many ranks wrote it: files with higher numbers are left-overs of an earlier dump with more
    # ranks under the same name, fewer files than that is a truncated dump
    with RestartFile(out[0]) as r:
        ranks = r.attribute("numRanks", np.int64)
    if ranks is not None and ranks.size:
        if len(out) < int(ranks[0]):
            raise B.SphError(f"restart: {path} was written by {int(ranks[0])} ranks, only {len(out)} files found")
        out = out[:int(ranks[0])]
    return out


def save(sim, path, iteration, num_particles_global=None, rank=0, num_ranks=1):
    """dump the owned particles of `sim` (a sim.HydroVeProp); returns the file written"""
    if num_particles_global is None:
        num_particles_global = sim.endIndex - sim.startIndex
    p = rank_path(path, rank, num_ranks)
    check(lib.sphb200_sim_save(sim.h, p.encode(), C.c_uint64(iteration), C.c_uint64(num_particles_global)))
    return p


class RestartFile:
    """one open restart file (IFileReader: stepAttribute / readField)"""

    def __init__(self, path):
        self.h = vp()
        check(lib.sphb200_restart_open(path.encode(), C.byref(self.h)))
        self.info = B.RestartInfo()
        try:
            check(lib.sphb200_restart_get_info(self.h, C.byref(self.info)))
        except Exception:
            self.close()
            raise

    def close(self):
        if self.h:
            lib.sphb200_restart_close(self.h)
            self.h = vp()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def attribute(self, name, dtype=np.float64):
        """values of a step attribute as a numpy array of `dtype` (float64, int64 or uint8), None if absent"""
        n = lib.sphb200_restart_attribute(self.h, name.encode(), None, 0)
        if n == -2:
            raise B.SphError(lib.sphb200_last_error().decode())
        if n < 0:
            return None
        out = np.zeros(max(n, 1), dtype)
        if lib.sphb200_restart_attribute(self.h, name.encode(), out.ctypes.data_as(vp), C.c_size_t(n)) == -2:
            raise B.SphError(lib.sphb200_last_error().decode())
        return out[:n]

    def read_fields(self, first, count, dtype, arrays, offset=0):
        """particles [first, first+count) of every field in `arrays` (name -> numpy array), written at `offset`"""
        f = B.Fields()
        for n in B.FIELD_NAMES:
            setattr(f, n, None)
        for n, a in arrays.items():
            setattr(f, n, a.ctypes.data + offset * a.itemsize)
        check(lib.sphb200_restart_read_fields(self.h, C.c_uint64(first), C.c_uint64(count), C.c_int(dtype),
                                              C.byref(f)))

    def has_field(self, name):
        return lib.sphb200_restart_read_bytes(self.h, name.encode(), C.c_uint64(0), C.c_uint64(0), None) == 0

    def read_bytes(self, name, first, count, out, offset=0):
        """particles [first, first+count) of a byte field ("rung") into out[offset:]"""
        dst = out[offset:offset + count]
        check(lib.sphb200_restart_read_bytes(self.h, name.encode(), C.c_uint64(first), C.c_uint64(count),
                                             dst.ctypes.data_as(vp)))

    def apply(self, sim):
        """time, time steps, constants, box and turbulence state -> sim"""
        check(lib.sphb200_restart_apply(self.h, sim.h))


def slice_for_rank(total, rank, num_ranks):
    """equal contiguous split of partitionRange (main/src/io/ifile_io_hdf5.cpp:131-148): the first total % P ranks get one more"""
    base, rem = divmod(total, num_ranks)
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


def load(sim, path, rank=0, num_ranks=1):
    """read this rank's share of a dump (written by any number of ranks) into `sim` and apply the step attributes;
    the caller runs the usual sync afterwards.  Returns (iteration, numParticlesGlobal)."""
    files = [RestartFile(p) for p in find_files(path)]
    try:
        counts = [int(f.info.numParticles) for f in files]
        total = sum(counts)
        n_glob = int(files[0].info.numParticlesGlobal)
        if total != n_glob:
            raise B.SphError(f"restart: files hold {total} particles, numParticlesGlobal says {n_glob}")
        if num_ranks == len(files):
            # same partition as the dump: every rank takes its own file back, nothing moves in the first sync
            first, count = sum(counts[:rank]), counts[rank]
        else:
            first, count = slice_for_rank(total, rank, num_ranks)
        real = sim.real
        arrays = {n: np.empty(count, np.float64 if n in B.ALWAYS_F64 else real) for n in B.CONSERVED}
        # "rung" is a conserved field of the block-time-step propagators only (ve_hydro_bdt.hpp:94)
        bdt = hasattr(sim, "uploadRungs") and all(f.has_field("rung") for f in files)
        rung = np.zeros(count, np.uint8) if bdt else None
        filled, start = 0, 0
        for f, c in zip(files, counts):
            lo, hi = max(first, start), min(first + count, start + c)
            if hi > lo:
                f.read_fields(lo - start, hi - lo, sim.dtype, arrays, offset=filled)
                if bdt:
                    f.read_bytes("rung", lo - start, hi - lo, rung, offset=filled)
                filled += hi - lo
            start += c
        assert filled == count
        sim.upload(arrays)
        if bdt:
            sim.uploadRungs(rung)
        files[0].apply(sim)
        return int(files[0].info.iteration), n_glob
    finally:
        for f in files:
            f.close()
