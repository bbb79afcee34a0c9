"""Row-sharded hierarchy build: the host-side driver of gmg_shard_* (include/gravomg_b200.h) [not in
reference; SURVEY.md 8(e)].

One process per GPU. The C library runs the PHASES of a level on the rank's own range of fine
vertices and exposes the replicated per-vertex buffers it writes; this module runs the phases in
order and all-gathers those buffers between them:

    TorchComm      torch.distributed (backend "nccl" over NVLink / NVSwitch on the GPU box, "gloo" for the
                   CPU tests of the exchange logic) -- in-place all_gather_into_tensor of the owned slices;
    LoopbackComm   several VIRTUAL ranks in one process on one GPU: the same phases run for every
                   virtual rank in lock-step and the gather is a set of device-to-device copies. This is
                   how the sharded path is parity-tested against the single-GPU build on one B200.

The data path has no reduction and no all-to-all: every exchange is an all-gather of contiguous owned
slices. Results equal the single-GPU build bit for bit; P comes back sharded by rows. With
build_galerkin the rows of P are gathered, every rank forms the outer vectors of A_c = P^T A P that it
owns, and A_c is gathered (replicated: it is the next level's A).
"""
import ctypes as C

import numpy as np

from . import capi
from .capi import (GmgError, HierarchyOptions, LevelInfo, _f64p, _p,
                   H_SAMPLES, H_PARENTS, H_COARSE_COLPTR, H_COARSE_ROWIDX, H_COARSE_VAL, H_COARSE_POINTS, H_TRIANGLES,
                   H_NORMALS, H_FAN_PTR, H_FAN, H_P_ROWPTR, H_P_COLIDX, H_P_VAL, H_AC_COLPTR, H_AC_ROWIDX, H_AC_VAL)

(PH_LEVEL_BEGIN, PH_SAMPLER_BUILD, PH_SAMPLER_ROUND, PH_SAMPLER_CHECK, PH_SAMPLES, PH_PARENTS_PULL, PH_PARENTS_CHECK,
 PH_PARENTS_PRED, PH_PARENTS_RESOLVE, PH_COARSE_ROWS, PH_COARSE_PACK, PH_COARSE_ASSEMBLE, PH_PROLONG,
 PH_LEVEL_END, PH_RADIUS, PH_P_COUNT, PH_P_PACK, PH_GALERKIN_ROWS, PH_GALERKIN_PACK, PH_GALERKIN_ASSEMBLE,
 PH_LEVEL_STATUS) = range(21)
(XB_STATE, XB_CHANGED, XB_DIST, XB_PRED, XB_PARENTS, XB_COARSE_COUNT, XB_ROWS_ID, XB_ROWS_VAL, XB_LEN_PARTIAL,
 XB_P_COUNT, XB_P_COL, XB_P_VAL, XB_STATUS) = range(13)
PHASE_NAMES = ["level_begin", "sampler_build", "sampler_round", "sampler_check", "samples", "parents_pull", "parents_check",
               "parents_pred", "parents_resolve", "coarse_rows", "coarse_pack", "coarse_assemble", "prolong", "level_end",
               "radius", "p_count", "p_pack", "galerkin_rows", "galerkin_pack", "galerkin_assemble", "level_status"]
V_BUFFERS = (XB_ROWS_ID, XB_ROWS_VAL, XB_P_COL, XB_P_VAL)      # slices of different sizes (packed rows)
BUFFER_NAMES = ["state", "changed", "dist", "pred", "parents", "coarse_count", "rows_id", "rows_val", "len_partial",
                "p_count", "p_col", "p_val", "status"]


def partition(count, world, rank):
    """Contiguous split used by the C library (shard.cu split()): slices of a multiple of 32 entries.
    Returns (slice_length, lo, hi)."""
    chunk = ((count + world - 1) // world + 31) // 32 * 32
    if chunk == 0:
        chunk = 32
    return chunk, min(count, rank * chunk), min(count, (rank + 1) * chunk)


class _DeviceBytes:
    """A raw device range as a __cuda_array_interface__ object, so torch can wrap it without copying."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


def device_bytes_as_tensor(ptr, nbytes, device):
    import torch
    return torch.as_tensor(_DeviceBytes(ptr, nbytes), device=device)


class TorchComm:
    """All-gather through torch.distributed. `buffers` are byte tensors that alias the library's
    exchange buffers (CUDA for nccl). One rank per process."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.bytes_gathered = 0

    def all_gather_inplace(self, full, slice_bytes):
        """full: uint8 tensor of world * slice_bytes; rank r's slice is [r*slice_bytes, (r+1)*slice_bytes)."""
        mine = full[self.rank * slice_bytes:(self.rank + 1) * slice_bytes]
        self.dist.all_gather_into_tensor(full, mine, group=self.group)
        self.bytes_gathered += int(full.numel())

    def all_gather_v(self, full, byte_offsets):
        """Slices of different sizes: rank r's slice is full[byte_offsets[r]:byte_offsets[r+1]] and already in
        place. The slices are nearly equal (contiguous ranges of rows), so they travel as ONE equal-size
        all_gather_into_tensor through a padded staging tensor and are copied into place; NCCL's list
        all_gather with uneven sizes (grouped broadcasts) measured 2.4x slower on 8 B200."""
        import torch
        sizes = [byte_offsets[r + 1] - byte_offsets[r] for r in range(self.world)]
        if len(set(sizes)) == 1 and int(full.numel()) == self.world * sizes[0]:
            self.dist.all_gather_into_tensor(full, full[byte_offsets[self.rank]:byte_offsets[self.rank + 1]], group=self.group)
        else:
            smax = (max(sizes) + 15) // 16 * 16
            staging = torch.empty(self.world * smax, dtype=torch.uint8, device=full.device)
            mine = staging[self.rank * smax:(self.rank + 1) * smax]
            mine[:sizes[self.rank]].copy_(full[byte_offsets[self.rank]:byte_offsets[self.rank + 1]])
            self.dist.all_gather_into_tensor(staging, mine, group=self.group)
            for r in range(self.world):
                if r != self.rank and sizes[r]:
                    full[byte_offsets[r]:byte_offsets[r + 1]].copy_(staging[r * smax:r * smax + sizes[r]])
        self.bytes_gathered += int(full.numel())


def upload_sharded(ctx, comm, points, colptr, rowidx, a_val=None, device=None):
    """Row-sharded upload of a level-0 problem every rank holds on the host: each rank copies only the slice
    of its own vertices over PCIe (host -> device) and the slices are all-gathered over NVLink, so the whole
    mesh crosses PCIe once per BOX instead of once per GPU. Only the owned slices of the host arrays are read.
    Returns (capi.Input, bytes this rank copied from the host)."""
    import torch
    n, nnz = int(colptr.shape[0]) - 1, int(rowidx.shape[0])
    world, rank = comm.world, comm.rank
    device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
    inp = capi.Input.device_empty(ctx, n, nnz, with_a=a_val is not None)
    bounds = [partition(n, world, r)[1] for r in range(world)] + [n]
    lo, hi = bounds[rank], bounds[rank + 1]
    h2d = 0

    def put(dst, src_np):
        nonlocal h2d
        t = torch.from_numpy(src_np)
        dst.copy_(t.view(torch.uint8) if t.dtype != torch.uint8 else t, non_blocking=True)
        h2d += src_np.nbytes

    # positions: x | y | z, each coordinate a gather of its own
    pts = device_bytes_as_tensor(inp.device_ptr(0), 24 * n, device)
    for k in range(3):
        col = np.ascontiguousarray(points[lo:hi, k])
        off = [8 * (k * n + b) for b in bounds]
        put(pts[off[rank]:off[rank + 1]], col.view(np.uint8))
        comm.all_gather_v(pts, off)
    # colptr: entries [lo, hi) per rank, the last rank also owns entry n
    cp = device_bytes_as_tensor(inp.device_ptr(1), 4 * (n + 1), device)
    off = [4 * b for b in bounds]
    off[-1] = 4 * (n + 1)
    last = hi + 1 if rank == world - 1 else hi
    put(cp[off[rank]:off[rank + 1]], np.ascontiguousarray(colptr[lo:last]).view(np.uint8))
    comm.all_gather_v(cp, off)
    # entries: slices start where the ranks' first columns start (host colptr is complete on every rank)
    e_off = [int(colptr[b]) for b in bounds]
    ri = device_bytes_as_tensor(inp.device_ptr(2), 4 * nnz, device)
    put(ri[4 * e_off[rank]:4 * e_off[rank + 1]], np.ascontiguousarray(rowidx[e_off[rank]:e_off[rank + 1]]).view(np.uint8))
    comm.all_gather_v(ri, [4 * e for e in e_off])
    if a_val is not None:
        av = device_bytes_as_tensor(inp.device_ptr(3), 8 * nnz, device)
        put(av[8 * e_off[rank]:8 * e_off[rank + 1]], np.ascontiguousarray(a_val[e_off[rank]:e_off[rank + 1]]).view(np.uint8))
        comm.all_gather_v(av, [8 * e for e in e_off])
    torch.cuda.current_stream().synchronize()       # the copies ran on torch's stream; commit runs on the context's
    inp.commit()
    inp.h2d_bytes = h2d
    return inp, h2d


class ShardRank:
    """One rank's gmg_shard handle."""

    def __init__(self, ctx, rank, world, inp, opt):
        self.ctx, self.rank, self.world, self._inp = ctx, rank, world, inp
        lib = ctx.lib
        lib.gmg_shard_phase.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int64)]
        lib.gmg_shard_buffer.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int32),
                                         C.POINTER(C.c_int64)]
        lib.gmg_shard_destroy.argtypes = [C.c_void_p]
        lib.gmg_shard_destroy.restype = None
        lib.gmg_shard_num_levels.argtypes = [C.c_void_p]
        lib.gmg_shard_level_info.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        lib.gmg_shard_fetch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        lib.gmg_shard_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]
        h = C.c_void_p()
        ctx._check(lib.gmg_shard_create(ctx.h, C.c_int(rank), C.c_int(world), inp.h, C.byref(opt), C.byref(h)))
        self.h = h
        self.io = (C.c_int64 * 8)()

    def phase(self, ph):
        self.ctx._check(self.ctx.lib.gmg_shard_phase(self.ctx.h, self.h, C.c_int(ph), self.io))
        return self.io

    def buffer_v(self, which):
        """(device pointer, byte offsets of the ranks' slices [world + 1]) of a variable-size exchange buffer."""
        lib = self.ctx.lib
        lib.gmg_shard_buffer_v.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int32),
                                           C.POINTER(C.c_int64)]
        ptr, eb = C.c_void_p(), C.c_int32()
        off = (C.c_int64 * (self.world + 1))()
        rc = lib.gmg_shard_buffer_v(self.h, C.c_int(which), C.byref(ptr), C.byref(eb), off)
        if rc != 0:
            raise GmgError(rc, "gmg_shard_buffer_v")
        return ptr.value, [int(o) * int(eb.value) for o in off]

    def buffer(self, which):
        """(device pointer, bytes per slice) of an exchange buffer."""
        ptr, eb, sl = C.c_void_p(), C.c_int32(), C.c_int64()
        rc = self.ctx.lib.gmg_shard_buffer(self.h, C.c_int(which), C.byref(ptr), C.byref(eb), C.byref(sl))
        if rc != 0:
            raise GmgError(rc, "gmg_shard_buffer")
        return ptr.value, int(eb.value) * int(sl.value)

    def close(self):
        if getattr(self, "h", None) and getattr(self.ctx, "h", None):
            self.ctx.lib.gmg_shard_destroy(self.h)
        self.h = None

    __del__ = close


class NativeComm:
    """The library's own NCCL communicator (csrc/shard_nccl.cu): gmg_shard_build_nccl runs the whole driver loop in
    C++ and issues ncclAllGather itself. The unique id travels through torch.distributed once, at set-up."""

    def __init__(self, ctx, group=None):
        import torch.distributed as dist
        lib = ctx.lib
        self.ctx = ctx
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        lib.gmg_nccl_unique_id.argtypes = [C.c_void_p]
        lib.gmg_nccl_init.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        lib.gmg_nccl_destroy.argtypes = [C.c_void_p]
        lib.gmg_nccl_destroy.restype = None
        lib.gmg_shard_build_nccl.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int64)]
        buf = C.create_string_buffer(128)
        if self.rank == 0:
            rc = lib.gmg_nccl_unique_id(buf)
            if rc != 0:
                raise GmgError(rc, "gmg_nccl_unique_id: libnccl.so.2 could not be opened")
        box = [bytes(buf.raw) if self.rank == 0 else None]
        dist.broadcast_object_list(box, src=0, group=group)
        self._id = C.create_string_buffer(box[0], 128)
        h = C.c_void_p()
        ctx._check(lib.gmg_nccl_init(ctx.h, self._id, self.rank, self.world, C.byref(h)))
        self.h = h

    def build(self, shard, build_galerkin):
        stats = (C.c_int64 * 2)()
        self.ctx._check(self.ctx.lib.gmg_shard_build_nccl(self.ctx.h, shard.h, self.h, 1 if build_galerkin else 0, stats))
        return int(stats[0]), int(stats[1])

    def close(self):
        if getattr(self, "h", None) and getattr(self.ctx, "h", None):
            self.ctx.lib.gmg_nccl_destroy(self.h)
        self.h = None

    __del__ = close


class ShardedHierarchy:
    """Row-sharded build. `ranks` is the list of ranks THIS process runs: one (its torch.distributed rank)
    in production, all of them as virtual ranks with comm=None (loop-back on one GPU) in the tests."""

    def __init__(self, ctx, inp, levels, world, ranks, comm=None, ratio=8.0,
                 weighting=capi.BARYCENTRIC, sampler_mode=capi.SAMPLER_HASHED, seed=0, radius_policy=0,
                 radius_override=None, device=None, profile=False, build_galerkin=False, native=None):
        import torch
        self.torch = torch
        self.ctx, self.world, self.comm = ctx, world, comm
        # The exchange runs on torch's current stream. If `ctx` was created on that very stream
        # (s = torch.cuda.Stream(); torch.cuda.set_stream(s); Context(device, s.cuda_stream)) the phases and
        # the exchanges are ordered by the stream itself; otherwise (the library's own stream, or torch's
        # legacy default stream 0, which the library never runs on) every exchange is fenced by a device
        # synchronise on both sides.
        self.same_stream = ctx.stream_handle != 0 and ctx.stream_handle == torch.cuda.current_stream().cuda_stream
        if comm is not None and len(ranks) != 1:
            raise ValueError("with a communicator every process runs exactly one rank")
        if comm is None and sorted(ranks) != list(range(world)):
            raise ValueError("the loop-back exchange needs every virtual rank in this process")
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        ro = None
        if radius_override is not None:
            ro = np.zeros(max(levels - 1, 1), dtype=np.float64)
            ro[:len(radius_override)] = radius_override
        self._ro = ro
        opt = HierarchyOptions(levels, ratio, weighting, sampler_mode, seed, radius_policy,
                               _p(ro, _f64p) if ro is not None else None, 1 if build_galerkin else 0)
        self.build_galerkin = bool(build_galerkin)
        # `inp` (capi.Input): the whole level-0 problem resident on this device; virtual ranks share it
        self.shards = [ShardRank(ctx, r, world, inp, opt) for r in ranks]
        self.exchanges = 0
        self.exchange_bytes = 0
        self.profile = {} if profile else None          # name -> [event pairs]; see phase_ms()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        if native is not None:                  # the C++ driver: one call, NCCL issued by the library itself
            if len(ranks) != 1 or profile:
                raise ValueError("the native driver runs one rank per process and has no per-phase profile")
            ex, by = native.build(self.shards[0], self.build_galerkin)
            self.exchanges, self.exchange_bytes = ex - getattr(native, "_seen_ex", 0), by - getattr(native, "_seen_by", 0)
            native._seen_ex, native._seen_by = ex, by
        else:
            self._build()
        ev1.record()
        torch.cuda.synchronize()
        self.build_ms = ev0.elapsed_time(ev1)
        self.num_levels = int(ctx.lib.gmg_shard_num_levels(self.shards[0].h))

    # ---- exchange ------------------------------------------------------------------------------------
    def _gather(self, which):
        return self._timed("gather_" + BUFFER_NAMES[which], lambda: self._gather_now(which))

    def _gather_v(self, which):
        t = self.torch
        bufs = [s.buffer_v(which) for s in self.shards]
        off = bufs[0][1]
        total = off[-1]
        self.exchanges += 1
        self.exchange_bytes += total
        if total == 0:
            return
        if not self.same_stream:
            t.cuda.synchronize()
        if self.comm is not None:
            self.comm.all_gather_v(device_bytes_as_tensor(bufs[0][0], total, self.device), off)
        else:
            views = [device_bytes_as_tensor(p, total, self.device) for p, _ in bufs]
            for src, shard in enumerate(self.shards):
                r = shard.rank
                if off[r + 1] > off[r]:
                    for dst in range(len(views)):
                        if dst != src:
                            views[dst][off[r]:off[r + 1]].copy_(views[src][off[r]:off[r + 1]])
        if not self.same_stream:
            t.cuda.synchronize()

    def _gather_now(self, which):
        if which in V_BUFFERS:
            return self._gather_v(which)
        t = self.torch
        bufs = [s.buffer(which) for s in self.shards]
        slice_bytes = bufs[0][1]
        total = slice_bytes * self.world
        self.exchanges += 1
        self.exchange_bytes += total
        if not self.same_stream:
            t.cuda.synchronize()
        if self.comm is not None:
            full = device_bytes_as_tensor(bufs[0][0], total, self.device)
            self.comm.all_gather_inplace(full, slice_bytes)
            if not self.same_stream:
                t.cuda.synchronize()
            return
        # loop-back: copy every virtual rank's owned slice into all the other replicas
        views = [device_bytes_as_tensor(p, total, self.device) for p, _ in bufs]
        for src, shard in enumerate(self.shards):
            r = shard.rank
            piece = views[src][r * slice_bytes:(r + 1) * slice_bytes]
            for dst in range(len(views)):
                if dst != src:
                    views[dst][r * slice_bytes:(r + 1) * slice_bytes].copy_(piece)
        if not self.same_stream:
            t.cuda.synchronize()

    def _timed(self, name, fn):
        if self.profile is None:
            return fn()
        t = self.torch
        e0, e1 = t.cuda.Event(enable_timing=True), t.cuda.Event(enable_timing=True)
        e0.record()
        out = fn()
        e1.record()
        self.profile.setdefault(name, []).append((e0, e1))
        return out

    def phase_ms(self):
        """Device milliseconds per phase / exchange name, summed over levels (profile=True builds only)."""
        self.torch.cuda.synchronize()
        return {k: float(sum(a.elapsed_time(b) for a, b in v)) for k, v in (self.profile or {}).items()}

    def _phase(self, ph):
        def run():
            return [s.phase(ph) for s in self.shards][0]
        return self._timed(PHASE_NAMES[ph], run)

    def _build(self):
        while True:
            if not self._phase(PH_LEVEL_BEGIN)[0]:
                break
            self._gather(XB_LEN_PARTIAL)
            self._phase(PH_RADIUS)
            self._phase(PH_SAMPLER_BUILD)
            self._gather(XB_STATE)
            while self._phase(PH_SAMPLER_CHECK)[0] > 0:
                self._phase(PH_SAMPLER_ROUND)
                self._gather(XB_STATE)
            self._phase(PH_SAMPLES)
            while True:
                self._phase(PH_PARENTS_PULL)
                self._gather(XB_DIST)
                self._gather(XB_CHANGED)
                if self._phase(PH_PARENTS_CHECK)[0] == 0:
                    break
            self._phase(PH_PARENTS_PRED)
            self._gather(XB_PRED)
            self._phase(PH_PARENTS_RESOLVE)
            self._gather(XB_PARENTS)
            self._phase(PH_COARSE_ROWS)
            self._gather(XB_COARSE_COUNT)
            self._phase(PH_COARSE_PACK)
            self._gather(XB_ROWS_ID)
            self._gather(XB_ROWS_VAL)
            self._phase(PH_COARSE_ASSEMBLE)
            self._phase(PH_PROLONG)
            if self.build_galerkin:
                self._phase(PH_P_COUNT)
                self._gather(XB_P_COUNT)
                self._phase(PH_P_PACK)
                self._gather(XB_P_COL)
                self._gather(XB_P_VAL)
                self._phase(PH_GALERKIN_ROWS)
                self._gather(XB_COARSE_COUNT)
                self._phase(PH_GALERKIN_PACK)
                self._gather(XB_ROWS_ID)
                self._gather(XB_ROWS_VAL)
                self._phase(PH_GALERKIN_ASSEMBLE)
            # the deferred checks of the level on the values of ALL ranks: every rank raises the same error together
            self._phase(PH_LEVEL_STATUS)
            self._gather(XB_STATUS)
            self._phase(PH_LEVEL_END)

    # ---- results -------------------------------------------------------------------------------------
    def info(self, level, shard=0):
        li = LevelInfo()
        lo, hi = C.c_int64(), C.c_int64()
        s = self.shards[shard]
        self.ctx._check(self.ctx.lib.gmg_shard_level_info(s.h, C.c_int(level), C.byref(li), C.byref(lo), C.byref(hi)))
        d = li.as_dict()
        d["row_lo"], d["row_hi"] = int(lo.value), int(hi.value)
        return d

    def fetch(self, level, which, shard=0, out=None):
        i = self.info(level, shard)
        n, m, T, own = i["n"], i["m"], i["T"], i["row_hi"] - i["row_lo"]
        count, dtype = {
            H_SAMPLES: (m, np.int64), H_PARENTS: (n, np.int64), H_COARSE_COLPTR: (m + 1, np.int32),
            H_COARSE_ROWIDX: (i["nnz_c"], np.int32), H_COARSE_VAL: (i["nnz_c"], np.float64),
            H_COARSE_POINTS: (3 * m, np.float64), H_TRIANGLES: (3 * T, np.int64), H_NORMALS: (3 * T, np.float64),
            H_FAN_PTR: (m + 1, np.int64), H_FAN: (3 * T, np.int64), H_P_ROWPTR: (own + 1, np.int32),
            H_P_COLIDX: (i["nnz_p"], np.int32), H_P_VAL: (i["nnz_p"], np.float64), H_AC_COLPTR: (m + 1, np.int32),
            H_AC_ROWIDX: (i["nnz_ac"], np.int32), H_AC_VAL: (i["nnz_ac"], np.float64)}[which]
        if out is None:
            out = np.empty(max(count, 1), dtype=dtype)
        elif out.dtype != dtype or out.ndim != 1 or out.shape[0] < count:
            raise ValueError("fetch: `out` must be a 1-D array of the array's dtype and length")
        s = self.shards[shard]
        self.ctx._check(self.ctx.lib.gmg_shard_fetch(self.ctx.h, s.h, C.c_int(level), C.c_int(which),
                                                     out.ctypes.data_as(C.c_void_p)))
        out = out[:count]
        if which == H_COARSE_POINTS:
            return out.reshape((m, 3), order="F")
        if which in (H_TRIANGLES, H_NORMALS):
            return out.reshape(T, 3)
        return out

    def level_arrays(self, level):
        """Like Hierarchy.level_arrays; the rows of P are concatenated over the ranks this process runs
        (all of them under the loop-back exchange, so the result is the whole of P)."""
        f = self.fetch
        out = dict(radius=self.info(level)["radius"], sel=f(level, H_SAMPLES), par=f(level, H_PARENTS),
                   ccp=f(level, H_COARSE_COLPTR), cri=f(level, H_COARSE_ROWIDX), cval=f(level, H_COARSE_VAL),
                   cpts=f(level, H_COARSE_POINTS), tri=f(level, H_TRIANGLES), nrm=f(level, H_NORMALS),
                   fptr=f(level, H_FAN_PTR), fan=f(level, H_FAN))
        rowptr, cols, vals, base = [np.zeros(1, dtype=np.int32)], [], [], 0
        for k in range(len(self.shards)):
            rp = f(level, H_P_ROWPTR, k)
            rowptr.append(rp[1:] + base)
            cols.append(f(level, H_P_COLIDX, k))
            vals.append(f(level, H_P_VAL, k))
            base += int(rp[-1])
        out["prp"] = np.concatenate(rowptr).astype(np.int32)
        out["pci"] = np.concatenate(cols)
        out["pva"] = np.concatenate(vals)
        return out

    def galerkin(self, level, shard=0):
        """A_c of a level (replicated) as scipy CSC."""
        import scipy.sparse as sp
        m = self.info(level, shard)["m"]
        return sp.csc_matrix((self.fetch(level, H_AC_VAL, shard), self.fetch(level, H_AC_ROWIDX, shard),
                              self.fetch(level, H_AC_COLPTR, shard)), shape=(m, m))

    def close(self):
        for s in getattr(self, "shards", []):
            s.close()
        self.shards = []


# ---- row-sharded smoothing / V-cycle -----------------------------------------------------------------
(SV_X, SV_B, SV_R, SV_COLOUR, SV_MASK) = range(5)
(HP_A, HP_R, HP_P) = range(3)
(SO_GS_COLOUR, SO_RESIDUAL, SO_RESTRICT, SO_PROLONG, SO_ZERO_X, SO_SUMSQ_R, SO_GS_ALL, SO_VCYCLE_ALL) = range(8)


def halo_counts(ghost, chunk, world):
    """Entries of an ascending ghost list per owning rank (owner of vertex j = j // chunk)."""
    import torch
    if ghost.numel() == 0:
        return [0] * world
    return torch.bincount(torch.div(ghost, chunk, rounding_mode="floor"), minlength=world).tolist()


class HaloPlan:
    """Who sends what for one halo pattern. Per local rank: `ghost` (ascending global indices this rank
    reads but does not own, hence grouped by owner), `recv_counts` per owner, `send_idx` (the owned
    indices the other ranks read, grouped by reader, ascending per reader) and `send_counts`."""

    def __init__(self, ghost, recv_counts, send_idx, send_counts):
        self.ghost, self.recv_counts, self.send_idx, self.send_counts = ghost, recv_counts, send_idx, send_counts
        self.volume = sum(sum(c) for c in send_counts)
        self.empty_everywhere = False      # set by the solver once every rank's volume is known

    def select(self, keep_ghost, keep_send):
        """Sub-plan over the entries flagged by the boolean tensors (same rule on both sides of a pair)."""
        import torch
        ghost, rc, sidx, sc = [], [], [], []
        for g, r_counts, kg in zip(self.ghost, self.recv_counts, keep_ghost):
            ghost.append(g[kg])
            rc.append([int(p.sum()) for p in torch.split(kg, r_counts)])
        for si, s_counts, ks in zip(self.send_idx, self.send_counts, keep_send):
            sidx.append(si[ks])
            sc.append([int(p.sum()) for p in torch.split(ks, s_counts)])
        return HaloPlan(ghost, rc, sidx, sc)


def plan_over_dist(dist, group, ghost, chunk, world):
    """HaloPlan of ONE rank per process: the ghost list goes to the owners (all_to_all_single of the counts,
    then of the indices), which answer nothing -- what they receive IS their send list."""
    import torch
    recv_counts = halo_counts(ghost, chunk, world)
    rc = torch.tensor(recv_counts, dtype=torch.int64, device=ghost.device)
    sc = torch.empty_like(rc)
    dist.all_to_all_single(sc, rc, group=group)
    send_counts = sc.tolist()
    send_idx = torch.empty(sum(send_counts), dtype=torch.int64, device=ghost.device)
    dist.all_to_all_single(send_idx, ghost, output_split_sizes=send_counts, input_split_sizes=recv_counts, group=group)
    return HaloPlan([ghost], [recv_counts], [send_idx], [send_counts])


def exchange_over_dist(dist, group, plan, vec):
    """Owner values -> ghost slots of `vec` (one rank per process)."""
    import torch
    packed = vec[plan.send_idx[0]]
    recv = torch.empty(sum(plan.recv_counts[0]), dtype=vec.dtype, device=vec.device)
    dist.all_to_all_single(recv, packed, output_split_sizes=plan.recv_counts[0], input_split_sizes=plan.send_counts[0],
                           group=group)
    if recv.numel():
        vec[plan.ghost[0]] = recv


class ShardedSolver:
    """V-cycle on A x = b over a ShardedHierarchy built with build_galerkin=True. Operators run on the
    owned rows (gmg_shard_solver_op); ghost entries move between ranks as packed all-to-all exchanges
    (torch.distributed all_to_all_single over NCCL, or the loop-back between virtual ranks)."""

    def __init__(self, sh, inp, replicate_below=2_000_000):
        """replicate_below: levels >= 1 with at most this many vertices, and every coarser one, are finished
        whole on every rank (gather the right-hand side once, no halo exchange below); 0 = only the coarsest."""
        import torch
        self.torch, self.sh, self.ctx, self._inp = torch, sh, sh.ctx, inp
        self.comm, self.world, self.device = sh.comm, sh.world, sh.device
        self.same_stream = sh.same_stream
        lib = self.ctx.lib
        lib.gmg_shard_solver_create.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_void_p)]
        lib.gmg_shard_solver_destroy.argtypes = [C.c_void_p]
        lib.gmg_shard_solver_destroy.restype = None
        lib.gmg_shard_solver_num_levels.argtypes = [C.c_void_p]
        lib.gmg_shard_solver_level.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int64)]
        lib.gmg_shard_solver_vector.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        lib.gmg_shard_solver_halo_mask.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int64)]
        lib.gmg_shard_solver_op.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double)]
        self.ranks = [s.rank for s in sh.shards]
        self._vec_cache = {}
        self.h = []
        for s in sh.shards:
            h = C.c_void_p()
            self.ctx._check(lib.gmg_shard_solver_create(self.ctx.h, s.h, inp.h, int(replicate_below), C.byref(h)))
            self.h.append(h)
        self.num_levels = int(lib.gmg_shard_solver_num_levels(self.h[0]))
        self.exchanges = 0
        self.exchange_bytes = 0
        self.geom = []        # per level: list over local ranks of dict(n, lo, hi, chunk, colours)
        io = (C.c_int64 * 8)()
        for l in range(self.num_levels):
            g = []
            for h in self.h:
                self.ctx._check(lib.gmg_shard_solver_level(h, C.c_int(l), io))
                g.append(dict(n=int(io[0]), lo=int(io[1]), hi=int(io[2]), chunk=int(io[3]), colours=int(io[4]),
                              replicated=bool(io[5])))
            self.geom.append(g)
        self._fence()
        # halo plans: per level the operator's ghosts split by colour, the restriction's and the prolongation's
        self.plan_a, self.plan_colour, self.plan_r, self.plan_p = [], [], [], []
        for l in range(self.num_levels):
            pa = self._plan(l, HP_A)
            self.plan_a.append(pa)
            colour = [self._vec(k, l, SV_COLOUR) for k in range(len(self.h))]
            per_colour = []
            for col in range(self.geom[l][0]["colours"]):
                per_colour.append(pa.select([colour[k][pa.ghost[k]] == col for k in range(len(self.h))],
                                            [colour[k][pa.send_idx[k]] == col for k in range(len(self.h))]))
            self.plan_colour.append(per_colour)
            if l + 1 < self.num_levels:
                self.plan_r.append(self._plan(l, HP_R))
                self.plan_p.append(self._plan(l, HP_P))
        # an exchange nobody has anything to send in is skipped by every rank (one all-reduce decides, here)
        plans = self.plan_a + [p for per in self.plan_colour for p in per] + self.plan_r + self.plan_p
        vol = torch.tensor([p.volume for p in plans], dtype=torch.int64, device=self.device)
        if self.comm is not None:
            self.comm.dist.all_reduce(vol, group=self.comm.group)
        for p, v in zip(plans, vol.tolist()):
            p.empty_everywhere = v == 0
        self.last_ms = 0.0

    # ---- plumbing --------------------------------------------------------------------------------------
    def _fence(self):
        if not self.same_stream:
            self.torch.cuda.synchronize()

    def _vec(self, k, level, which):
        key = (k, level, which)
        hit = self._vec_cache.get(key)
        if hit is None:
            hit = self._vec_cache[key] = self._vec_wrap(k, level, which)
        return hit

    def _vec_wrap(self, k, level, which):
        """A solver vector as a torch tensor aliasing the library's device memory (stable for the solver's life)."""
        t = self.torch
        ptr = C.c_void_p()
        rc = self.ctx.lib.gmg_shard_solver_vector(self.h[k], C.c_int(level), C.c_int(which), C.byref(ptr))
        if rc != 0:
            raise GmgError(rc, "gmg_shard_solver_vector")
        g = self.geom[level][k]
        n_pad = max(g["n"], g["chunk"] * self.world)
        if which == SV_COLOUR:
            return device_bytes_as_tensor(ptr.value, 4 * g["n"], self.device).view(t.int32)
        if which == SV_MASK:
            return device_bytes_as_tensor(ptr.value, n_pad, self.device)
        return device_bytes_as_tensor(ptr.value, 8 * n_pad, self.device).view(t.float64)

    def _op(self, level, op, arg=0):
        out = C.c_double(0.0)
        vals = []
        for h in self.h:
            self.ctx._check(self.ctx.lib.gmg_shard_solver_op(self.ctx.h, h, C.c_int(level), C.c_int(op), C.c_int(arg), C.byref(out)))
            vals.append(out.value)
        return vals

    def _plan(self, level, pattern):
        """Ghost lists from the library's halo mask, then the matching send lists from the other ranks."""
        t = self.torch
        space = C.c_int64()
        target = level + 1 if pattern == HP_P else level
        chunk = self.geom[target][0]["chunk"]
        ghosts = []
        for k, h in enumerate(self.h):
            self.ctx._check(self.ctx.lib.gmg_shard_solver_halo_mask(self.ctx.h, h, C.c_int(level), C.c_int(pattern), C.byref(space)))
            self._fence()
            mask = self._vec(k, target, SV_MASK)[:int(space.value)]
            ghosts.append(t.nonzero(mask).flatten())
        if self.comm is not None:
            return plan_over_dist(self.comm.dist, self.comm.group, ghosts[0], chunk, self.world)
        recv_counts = [halo_counts(g, chunk, self.world) for g in ghosts]
        send_idx, send_counts = [], []
        for r in self.ranks:
            parts = [g[t.div(g, chunk, rounding_mode="floor") == r] for g in ghosts]
            send_idx.append(t.cat(parts))
            send_counts.append([int(p.numel()) for p in parts])
        return HaloPlan(ghosts, recv_counts, send_idx, send_counts)

    def _exchange(self, plan, level, which):
        """Owner values -> ghost slots of vector `which` of `level`, for the entries of `plan`."""
        t = self.torch
        if plan.empty_everywhere:
            return
        vecs = [self._vec(k, level, which) for k in range(len(self.h))]
        self._fence()
        self.exchanges += 1
        self.exchange_bytes += 8 * plan.volume
        if self.comm is not None:
            exchange_over_dist(self.comm.dist, self.comm.group, plan, vecs[0])
        else:
            packed = [v[idx] for v, idx in zip(vecs, plan.send_idx)]
            offs = [np.concatenate([[0], np.cumsum(c)]) for c in plan.send_counts]
            for q in range(len(self.h)):               # reader q receives, from every owner r, what r packed for q
                parts = [packed[r][int(offs[r][q]):int(offs[r][q + 1])] for r in range(len(self.h))]
                if plan.ghost[q].numel():
                    vecs[q][plan.ghost[q]] = t.cat(parts)
        self._fence()

    # ---- the cycle -------------------------------------------------------------------------------------
    def _smooth(self, level, sweeps, reverse):
        nc = self.geom[level][0]["colours"]
        for _ in range(sweeps):
            for kk in range(nc):
                col = nc - 1 - kk if reverse else kk
                self._op(level, SO_GS_COLOUR, col)
                self._exchange(self.plan_colour[level][col], level, SV_X)

    def _gather_vector(self, level, which):
        """All-gather the owned slices of a full-length vector (slice = the level's split, rank r owns slice r)."""
        t = self.torch
        chunk = self.geom[level][0]["chunk"]
        vecs = [self._vec(k, level, which)[:chunk * self.world] for k in range(len(self.h))]
        self._fence()
        self.exchanges += 1
        self.exchange_bytes += 8 * chunk * self.world
        if self.comm is not None:
            self.comm.all_gather_inplace(vecs[0].view(t.uint8), 8 * chunk)
        else:
            for src, r in enumerate(self.ranks):
                for dst in range(len(vecs)):
                    if dst != src:
                        vecs[dst][r * chunk:(r + 1) * chunk].copy_(vecs[src][r * chunk:(r + 1) * chunk])
        self._fence()

    def _vcycle(self, level, pre, post, coarse_sweeps):
        if self.geom[level][0]["replicated"]:
            # a small level (at least the coarsest): every rank gathers the whole right-hand side and finishes the
            # cycle redundantly (same kernels, same rows, same order as the single-GPU cycle) instead of exchanging
            # halos colour by colour
            self._gather_vector(level, SV_B)
            self._op(level, SO_VCYCLE_ALL, pre | (post << 8) | (coarse_sweeps << 16))
            return
        self._smooth(level, pre, False)
        self._op(level, SO_RESIDUAL)
        self._exchange(self.plan_r[level], level, SV_R)
        self._op(level, SO_RESTRICT)
        self._op(level + 1, SO_ZERO_X)
        self._vcycle(level + 1, pre, post, coarse_sweeps)
        if not self.geom[level + 1][0]["replicated"]:    # a replicated level's x is already whole on every rank
            self._exchange(self.plan_p[level], level + 1, SV_X)
        self._op(level, SO_PROLONG)
        self._exchange(self.plan_a[level], level, SV_X)
        self._smooth(level, post, True)

    def _residual_norm(self):
        t = self.torch
        self._op(0, SO_RESIDUAL)
        total = float(sum(self._op(0, SO_SUMSQ_R)))
        if self.comm is not None:
            v = t.tensor([total], dtype=t.float64, device=self.device)
            self.comm.dist.all_reduce(v, group=self.comm.group)
            total = float(v.item())
        return float(np.sqrt(total))

    def rhs_from_solution(self, xstar):
        """b = A x* on the owned rows, formed on the device (xstar: host array or device tensor, the WHOLE vector).
        Use vcycle(None, ...) afterwards."""
        t = self.torch
        n = self.geom[0][0]["n"]
        src = xstar if isinstance(xstar, t.Tensor) else t.from_numpy(np.ascontiguousarray(xstar, dtype=np.float64))
        self._fence()
        for k in range(len(self.h)):
            self._vec(k, 0, SV_X)[:n].copy_(src)
            self._vec(k, 0, SV_B).zero_()
        self._fence()
        self._op(0, SO_RESIDUAL)                       # r = 0 - A x*
        self._fence()
        for k in range(len(self.h)):
            g = self.geom[0][k]
            self._vec(k, 0, SV_B)[g["lo"]:g["hi"]] = -self._vec(k, 0, SV_R)[g["lo"]:g["hi"]]
        self._fence()

    def vcycle(self, b, x0=None, cycles=1, pre=2, post=2, coarse_sweeps=50, fetch=True):
        """b, x0: the WHOLE right-hand side / start vector (every rank passes the same); b = None keeps the
        right-hand side already on the device (rhs_from_solution). Returns (x, residuals): x holds the owned
        entries of the ranks this process runs (everything under the loop-back), NaN elsewhere (None if not fetch)."""
        t = self.torch
        n = self.geom[0][0]["n"]
        self._fence()
        for k in range(len(self.h)):
            if b is not None:
                self._vec(k, 0, SV_B)[:n].copy_(t.from_numpy(np.ascontiguousarray(b, dtype=np.float64)))
            if x0 is None:
                self._vec(k, 0, SV_X).zero_()
            else:
                self._vec(k, 0, SV_X)[:n].copy_(t.from_numpy(np.ascontiguousarray(x0, dtype=np.float64)))
        self._fence()
        res = [self._residual_norm()]
        e0, e1 = t.cuda.Event(enable_timing=True), t.cuda.Event(enable_timing=True)
        self.last_ms = 0.0
        for _ in range(cycles):
            self._fence()
            e0.record()
            self._vcycle(0, pre, post, coarse_sweeps)
            e1.record()
            t.cuda.synchronize()
            self.last_ms += e0.elapsed_time(e1)
            res.append(self._residual_norm())
        if not fetch:
            return None, np.array(res)
        x = np.full(n, np.nan)
        for k in range(len(self.h)):
            g = self.geom[0][k]
            x[g["lo"]:g["hi"]] = self._vec(k, 0, SV_X)[g["lo"]:g["hi"]].cpu().numpy()
        return x, np.array(res)

    def close(self):
        for h in getattr(self, "h", []):
            if getattr(self.ctx, "h", None):
                self.ctx.lib.gmg_shard_solver_destroy(h)
        self.h = []
