"""Row-slab decomposition of one large image across the GPUs of a box (BASELINE.json configs[4]).

One process per GPU (``torch.distributed``; NCCL on the box, gloo in the CPU tests).  Rank ``r`` owns the rows
``[r0, r1)`` of the ``Hg x W`` image, of the data and of BOTH planes of the dual variable; its local arrays carry
``halo_up`` rows above and ``halo_dn`` rows below (none at the image's own top / bottom edge).  Every operator on the
Condat-Vu TV path is a stencil of bounded reach (``Gradient``: 1 row, ``A^T A`` for a ``k``-row PSF: ``k - 1`` rows), so
one update of the owned rows needs ``x`` on 14 rows above / 15 below and ``z`` on 1 row above / below.

Per iteration NOTHING runs on the host but the one kernel launch: the fused kernel (``csrc/pcs_pds.cu``) stores the
boundary rows of ``x', z'`` straight into the neighbours' halo rows through NVLink (peer mappings from
``csrc/pcs_peer.cu``) and hands over through one counter word per neighbour: every block that owns rows of the 16-row
edge zone adds its share with a system-scope release as soon as it is done, and the neighbour's edge blocks of the next
step acquire the full count (``PCS_SLAB_TIMEOUT_S``, default 20 s, bounds the device-side wait).  The four norms of the stop test
are summed over ranks with one 32-byte all-reduce, only on the iterations whose stop test is live.

What runs once per solve (scattering ``y``'s halo rows to form ``A^T y``, filling the halos of ``x0, z0``) uses plain
``torch.distributed`` point-to-point calls (``gather_halo``), which is also what the gloo tests exercise.

Usage (every rank)::

    comm = SlabComm()                                   # default process group
    lay  = comm.layout(Hg, W)                           # this rank's rows: lay.r0, lay.r1
    F = (1/2) * SquaredL2Loss(dim=Hg*W, data=lay.local(y[lay.r0:lay.r1])) * Convolve2D(size=Hg*W, filter=psf, shape=(Hg, W))
    pds = PDS(dim=Hg*W, F=F, G=NonNegativeOrthant(Hg*W), H=lam*L21Norm(...), K=Gradient((Hg, W), kind='forward'),
              slab=comm, ...)
    est, _, diag = pds.iterate()                        # est[...] hold this rank's rows (z: [plane0 rows ; plane1 rows])

Operators and functionals describe the GLOBAL problem; data, ``x0``/``z0`` and the results are this rank's rows.
"""
from __future__ import annotations

import ctypes as C
import os
import weakref

import numpy as np

import torch
import torch.distributed as dist

HALO_UP = 14      # rows of x needed above an updated row: reach of A^T A for a 15-row PSF
HALO_DN = 15      # ... below: the same, plus the row whose primal update feeds the forward difference of the last row
Y_UP, Y_DN = 7, 8  # rows of y needed to form A^T y on the rows [r0, r1]


# ------------------------------------------------------------------------------------------------ host logic
class SlabLayout:
    """Row ranges and local-array geometry of one rank (pure host arithmetic; no CUDA, no communication)."""

    def __init__(self, Hg: int, W: int, world: int, rank: int, halo_up: int = HALO_UP, halo_dn: int = HALO_DN):
        if not (0 <= rank < world):
            raise ValueError('rank outside the group')
        if Hg < world * 2 * 16:
            raise ValueError(f'image of {Hg} rows is too small for {world} slabs (each needs >= 32 rows)')
        self.Hg, self.W, self.world, self.rank = int(Hg), int(W), int(world), int(rank)
        self.halo_up, self.halo_dn = int(halo_up), int(halo_dn)
        self.r0, self.r1 = self.rows_of(rank)
        self.n_own = self.r1 - self.r0
        self.ht = 0 if rank == 0 else self.halo_up             # halo rows present above / below the owned rows
        self.hb = 0 if rank == world - 1 else self.halo_dn
        self.H_loc = self.ht + self.n_own + self.hb
        self.grow0 = self.r0 - self.ht                         # global row of local row 0
        self.row_begin, self.row_end = self.ht, self.ht + self.n_own

    def rows_of(self, rank: int):
        """Balanced contiguous partition of the rows (the first ``Hg % world`` ranks get one row more)."""
        q, rem = divmod(self.Hg, self.world)
        r0 = rank * q + min(rank, rem)
        return r0, r0 + q + (1 if rank < rem else 0)

    def neighbour(self, side: int):
        """Layout of the rank above (``side = -1``) or below (``+1``), or None at the image edge."""
        r = self.rank + side
        if r < 0 or r >= self.world:
            return None
        return SlabLayout(self.Hg, self.W, self.world, r, self.halo_up, self.halo_dn)

    @property
    def plane(self) -> int:
        return self.H_loc * self.W

    def local(self, rows):
        """Wrap this rank's rows of an image-sized vector for use as functional data (``SquaredL2Loss(N, data=...)``)."""
        from .._array import SlabRows
        return SlabRows(rows, self.Hg * self.W)

    def own(self, t: torch.Tensor) -> torch.Tensor:
        """Owned rows (n_own x W view) of a local H_loc x W array."""
        return t.view(self.H_loc, self.W)[self.row_begin:self.row_end]

    def push_counts(self):
        """(nx_up, nx_dn, nz_up, nz_dn): how many of my first / last owned rows the neighbours hold as halo."""
        up, dn = self.neighbour(-1), self.neighbour(+1)
        return (up.hb if up else 0, dn.ht if dn else 0, 1 if up else 0, 1 if dn else 0)


def _stage_on_host(t: torch.Tensor, group=None) -> bool:
    """CUDA tensors travel through host memory when the group's backend is gloo (tests on one GPU, CPU runs)."""
    return t.is_cuda and dist.get_backend(group) == 'gloo'


def gather_halo(own: torch.Tensor, above: int, below: int, rank: int, world: int, group=None) -> torch.Tensor:
    """Rows ``[r0 - above, r1 + below)`` (clipped to the image) from this rank's rows ``own`` (n_own x W).

    Every rank calls it with the same ``above``/``below``; rank ``r`` sends its first ``below`` rows up and its last
    ``above`` rows down.  Works on any backend (CPU tensors under gloo, CUDA tensors under NCCL).
    """
    if _stage_on_host(own, group):
        return gather_halo(own.cpu(), above, below, rank, world, group).to(own.device)
    n, W = own.shape
    if above > n or below > n:
        raise ValueError('halo deeper than a slab')
    up = rank - 1 if rank > 0 else None
    dn = rank + 1 if rank < world - 1 else None
    g = (lambda r: r) if group is None else (lambda r: dist.get_global_rank(group, r))
    top = own.new_empty((above if up is not None else 0, W))
    bot = own.new_empty((below if dn is not None else 0, W))
    ops, keep = [], []
    if up is not None:
        if below:
            s = own[:below].contiguous()
            keep.append(s)
            ops.append(dist.P2POp(dist.isend, s, g(up), group))
        if above:
            ops.append(dist.P2POp(dist.irecv, top, g(up), group))
    if dn is not None:
        if above:
            s = own[n - above:].contiguous()
            keep.append(s)
            ops.append(dist.P2POp(dist.isend, s, g(dn), group))
        if below:
            ops.append(dist.P2POp(dist.irecv, bot, g(dn), group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return torch.cat([top, own, bot], dim=0)


class SlabComm:
    """The process group a slab-decomposed solve runs on (default: the world group)."""

    def __init__(self, group=None, lockstep=None):
        """``lockstep``: complete every update on all ranks before any rank launches the next one (a device
        synchronisation and a barrier per update).  The update kernels of neighbouring ranks hand over through flag
        words and SPIN until the neighbour's previous update has been published, which is only safe when the ranks
        really run concurrently, one per GPU.  Ranks that share a GPU (the single-GPU test set-up: gloo control plane,
        CUDA IPC between processes on one device) are time-sliced, so a spinning kernel could starve the kernel it
        waits for; in lockstep no kernel ever has to wait.  Default: on exactly when the backend is gloo."""
        if not dist.is_initialized():
            raise RuntimeError('SlabComm needs an initialised torch.distributed process group')
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.lockstep = (dist.get_backend(group) == 'gloo') if lockstep is None else bool(lockstep)

    def layout(self, Hg: int, W: int) -> SlabLayout:
        return SlabLayout(Hg, W, self.world, self.rank)

    def all_gather_object(self, obj):
        out = [None] * self.world
        dist.all_gather_object(out, obj, group=self.group)
        return out

    def barrier(self):
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        dist.barrier(group=self.group)

    def all_reduce_sum(self, t: torch.Tensor):
        if self.world > 1:
            if _stage_on_host(t, self.group):
                h = t.cpu()
                dist.all_reduce(h, op=dist.ReduceOp.SUM, group=self.group)
                t.copy_(h)
            else:
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t


# ------------------------------------------------------------------------------------------------ peer memory
class _RawCuda:
    """``__cuda_array_interface__`` holder: lets torch view memory this package allocated (csrc/pcs_peer.cu)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {'shape': (nbytes,), 'typestr': '|u1', 'data': (ptr, False), 'version': 2}


class PeerArena:
    """One exportable device allocation per rank, mapped by its row neighbours (CUDA IPC over NVLink)."""

    def __init__(self, comm: SlabComm, nbytes: int, meta: dict):
        from .. import _lib
        self._lib = _lib
        self.comm = comm
        self.nbytes = int(nbytes)
        hb = int(_lib.call_plain('pcs_peer_handle_bytes'))
        handle = C.create_string_buffer(hb)
        p = C.c_void_p()
        _lib.check(_lib.call_plain('pcs_peer_alloc', C.byref(p), self.nbytes, handle))
        self.ptr = int(p.value)
        self._raw = _RawCuda(self.ptr, self.nbytes)
        self.bytes = torch.as_tensor(self._raw, device=torch.device('cuda', torch.cuda.current_device()))
        infos = comm.all_gather_object({'handle': bytes(handle.raw), 'meta': meta})
        self.meta = [i['meta'] for i in infos]
        self.remote = {}
        for side in (-1, +1):
            r = comm.rank + side
            if 0 <= r < comm.world:
                q = C.c_void_p()
                _lib.check(_lib.call_plain('pcs_peer_open', infos[r]['handle'], C.byref(q)))
                self.remote[side] = int(q.value)
        comm.barrier()

    def view(self, offset: int, numel: int, dtype: torch.dtype) -> torch.Tensor:
        isz = torch.empty(0, dtype=dtype).element_size()
        return self.bytes[offset:offset + numel * isz].view(dtype)

    def close(self):
        if getattr(self, 'ptr', 0):
            torch.cuda.synchronize()
            self.comm.barrier()                          # nobody may still be storing into a neighbour
            for q in self.remote.values():
                self._lib.call_plain('pcs_peer_close', q)
            self.remote = {}
            self.comm.barrier()
            self.bytes = None
            self._lib.call_plain('pcs_peer_free', self.ptr)
            self.ptr = 0

    def __del__(self):
        try:
            if getattr(self, 'ptr', 0) and not self.remote:
                self._lib.call_plain('pcs_peer_free', self.ptr)
        except Exception:
            pass


_ARENAS = {}


def shared_arena(comm: SlabComm, nbytes: int, meta: dict) -> PeerArena:
    """Arena of this size for this group, created on first use and kept: allocation, IPC export and the neighbour
    mappings are collective and cost milliseconds, and successive solves of the same problem size reuse them (the flag
    words and the epoch counter live with the arena)."""
    key = (id(comm.group), comm.rank, comm.world, int(nbytes), repr(sorted(meta.items())))
    a = _ARENAS.get(key)
    if a is None or not a.ptr:
        a = _ARENAS[key] = PeerArena(comm, nbytes, meta)
        a.epoch = 0
    return a


def release_arenas():
    """Free every pooled arena (collective: every rank must call it)."""
    for a in list(_ARENAS.values()):
        a.close()
    _ARENAS.clear()


def _up256(n: int) -> int:
    return (n + 255) // 256 * 256


# ------------------------------------------------------------------------------------------------ fused plan
def make_slab_plan(solver, comm: SlabComm):
    """Slab-decomposed fused plan for a ``PrimalDualSplitting`` instance; raises if the problem is not supported."""
    from . import _fused as Fz
    if not solver._H or solver._dtype not in (torch.float32, torch.float64):
        raise NotImplementedError('slab decomposition needs a real-valued problem with a dual term')
    try:
        Hg, W, steps = Fz._match_K(solver.K, solver.dim)
        h = Fz._match_H(solver.H, Hg * W)
        g = Fz._match_G(solver.G, float(solver.tau))
        data_mode, y, A = Fz._match_F(solver.F, (Hg, W))
    except Fz._NoMatch as e:
        raise NotImplementedError(f'slab decomposition supports the fused total-variation problems only ({e})')
    if data_mode == Fz.DM_GIVEN:
        raise NotImplementedError('slab decomposition needs F in {None, 1/2 SquaredL2Loss, 1/2 SquaredL2Loss * separable '
                                  'Convolve2D}')
    return SlabPdsTvPlan(solver, comm, Hg, W, steps, data_mode, y, A, g, h)


class SlabPdsTvPlan:
    """State of a slab-decomposed Condat-Vu TV solve: peer arena, local geometry, per-step peer descriptor."""

    def __init__(self, solver, comm, Hg, W, steps, data_mode, y, A, g, h):
        from . import _fused as Fz
        from .. import _kernels as K
        from .. import _lib
        from .._array import to_device
        self.solver, self.comm = weakref.proxy(solver), comm
        self.lay = lay = comm.layout(Hg, W)
        self.steps, self.data_mode = steps, data_mode
        self.g_mode, self.g_a, self.g_b = g
        self.h_mode, self.lam = h
        self.dtype = dt = solver._dtype
        self.sfx = '_f32' if dt == torch.float32 else '_f64'
        isz = 4 if dt == torch.float32 else 8
        dev = torch.device('cuda', torch.cuda.current_device())
        # ---- peer arena: [x0 | x1 | z0 | z1 | flags] ; z buffers hold two H_loc x W planes
        nx, nz = lay.plane, 2 * lay.plane
        self.off_x = [0, _up256(nx * isz)]
        self.off_z = [self.off_x[1] + _up256(nx * isz), 0]
        self.off_z[1] = self.off_z[0] + _up256(nz * isz)
        self.off_flags = self.off_z[1] + _up256(nz * isz)
        total = self.off_flags + 256
        meta = {'off_x': self.off_x, 'off_z': self.off_z, 'off_flags': self.off_flags, 'H_loc': lay.H_loc,
                'grow0': lay.grow0}
        self.arena = shared_arena(comm, total, meta)
        self.xbuf = [self.arena.view(o, nx, dt) for o in self.off_x]
        self.zbuf = [self.arena.view(o, nz, dt) for o in self.off_z]
        # flag words (uint32): [0] written by the rank above, [1] by the rank below
        self.flags = self.arena.view(self.off_flags, 64, torch.int32)
        # ---- filter taps, corner corrections, data term (local rows [row_begin, row_end] of A^T y)
        self.ntaps, self.crow, self.ccol, self.edge = 1, None, None, None
        self.nl = self.nr = self.nt = self.nb = 0
        self.b = None
        if data_mode != Fz.DM_NONE:
            from .._array import SlabRows
            if not isinstance(y, SlabRows):
                raise ValueError('slab mode: pass this rank\'s rows of the data as SlabLayout.local(rows)')
            y_own = to_device(y.local, dt)
            if y_own.numel() != lay.n_own * W:
                raise ValueError(f'slab mode: the data must hold this rank\'s {lay.n_own} rows ({lay.n_own * W} values), '
                                 f'got {y_own.numel()}')
            self._y_own = y_own.view(lay.n_own, W)
            self._A = A
            self.b = torch.zeros(lay.plane, dtype=dt, device=dev)
            if data_mode == Fz.DM_CONV:
                ccol, Et, Eb = Fz.composite_filter(A.hcol, A.offset[0])
                crow, El, Er = Fz.composite_filter(A.hrow, A.offset[1])
                n = max(ccol.size, crow.size)
                if (n - 1) // 2 > HALO_UP:
                    raise NotImplementedError('PSF taller than the slab halo')

                def centred(c):
                    out = np.zeros(n)
                    p = (n - c.size) // 2
                    out[p:p + c.size] = c
                    return out
                self.ntaps = n
                self.crow, self.ccol = np.ascontiguousarray(centred(crow)), np.ascontiguousarray(centred(ccol))
                pack = np.zeros((4, Fz.MAX_HL, Fz.MAX_HL))
                for k, E in enumerate((El, Er, Et, Eb)):
                    pack[k, :E.shape[0], :E.shape[1]] = E
                self.nl, self.nr, self.nt, self.nb = El.shape[0], Er.shape[0], Et.shape[0], Eb.shape[0]
                self.edge = to_device(pack.reshape(-1), dt)
            self._b_ready = False
        nbytes = int(_lib.call_plain('pcs_pds_tv_workspace_bytes'))
        self.ws = torch.zeros(nbytes // 8, dtype=torch.float64, device=dev)
        self.err = torch.zeros(4, dtype=torch.int32, device=dev)
        self._desc = np.zeros(20, dtype=np.int64)
        self._K = K
        self._static_desc()
        comm.barrier()

    def _exchange_rows(self, buf, above, below):
        """Fill the first ``above`` / last ``below`` rows of ``buf`` ((above + n_own + below) x W, own rows already in the
        middle) with the neighbours' rows; rows beyond the image edge are left untouched (callers keep them zero)."""
        comm, lay = self.comm, self.lay
        rank, world, n = comm.rank, comm.world, lay.n_own
        g = (lambda r: r) if comm.group is None else (lambda r: dist.get_global_rank(comm.group, r))
        if _stage_on_host(buf, comm.group):               # gloo control plane (single-GPU tests): through the host
            ext = gather_halo(buf[above:above + n].cpu(), above, below, rank, world, comm.group)
            lo = above if rank == 0 else 0
            buf[lo:lo + ext.shape[0]].copy_(ext)
            return
        ops = []
        if rank > 0:
            if below:
                ops.append(dist.P2POp(dist.isend, buf[above:above + below], g(rank - 1), comm.group))
            if above:
                ops.append(dist.P2POp(dist.irecv, buf[:above], g(rank - 1), comm.group))
        if rank < world - 1:
            if above:
                ops.append(dist.P2POp(dist.isend, buf[n:n + above], g(rank + 1), comm.group))
            if below:
                ops.append(dist.P2POp(dist.irecv, buf[above + n:], g(rank + 1), comm.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()

    def begin_solve(self):
        """Per-solve precomputation (collective: one halo exchange of ``y``): the data term on the local rows
        ``[row_begin, row_end]`` -- ``y`` itself, or ``A^T y`` for the deconvolution problems.  The extended copy of
        ``y`` and the local convolution operator are set up once per plan; every solve re-reads ``y`` (it may have been
        modified in place), exchanges its boundary rows and re-forms ``A^T y``."""
        from . import _fused as Fz
        if self.data_mode == Fz.DM_NONE:
            return
        lay, comm, W = self.lay, self.comm, self.lay.W
        bv = self.b.view(lay.H_loc, W)
        above, below = (0, 1) if self.data_mode == Fz.DM_IDENTITY else (Y_UP, Y_DN)
        if getattr(self, '_ext', None) is None:
            self._ext = torch.zeros(above + lay.n_own + below, W, dtype=self.dtype, device=self.b.device)
            if self.data_mode == Fz.DM_CONV:
                # rows beyond the image edge stay zero = the zero boundary of the convolution; the first / last rank
                # simply drop them from the local operator's array
                from ..linop.conv import Convolve2D
                self._ext_lo = above if comm.rank == 0 else 0
                self._ext_hi = above + lay.n_own + (0 if comm.rank == comm.world - 1 else below)
                rows = self._ext_hi - self._ext_lo
                self._A_loc = Convolve2D(size=rows * W, filter=self._A.filter, shape=(rows, W))
        ext = self._ext
        ext[above:above + lay.n_own].copy_(self._y_own)
        self._exchange_rows(ext, above, below)
        if self.data_mode == Fz.DM_IDENTITY:
            keep = lay.n_own + (1 if comm.rank < comm.world - 1 else 0)
            bv[lay.row_begin:lay.row_begin + keep] = ext[:keep]
        else:
            # A^T y on the rows [r0, r1]: y with Y_UP rows above / Y_DN below, filtered with the zero boundary (the
            # true one at the image edge), keeping the rows at least Y_UP away from a cut
            loc = ext[self._ext_lo:self._ext_hi]
            bt = self._A_loc._adjoint(loc.reshape(-1)).view(loc.shape)
            skip = Y_UP if comm.rank > 0 else 0
            keep = lay.n_own + (1 if comm.rank < comm.world - 1 else 0)
            bv[lay.row_begin:lay.row_begin + keep] = bt[skip:skip + keep]
        self._b_ready = True

    # -- peer descriptor (include/pcsb200.h: pcs_pds_tv_step_slab) --------------------------------------------------
    def _static_desc(self):
        lay, d, m = self.lay, self._desc, self.arena.meta
        nx_up, nx_dn, nz_up, nz_dn = lay.push_counts()
        d[8], d[9], d[10], d[11] = nx_up, nx_dn, nz_up, nz_dn
        for side, (ip, idr, iw, isg) in ((-1, (4, 6, 12, 14)), (+1, (5, 7, 13, 15))):
            r = self.comm.rank + side
            if side in self.arena.remote:
                d[ip] = m[r]['H_loc'] * lay.W
                d[idr] = lay.grow0 - m[r]['grow0']
                d[iw] = self.flags.data_ptr() + (0 if side < 0 else 4)
                # the neighbour waits on the word reserved for ITS neighbour on my side: I am "below" the rank above
                d[isg] = self.arena.remote[side] + m[r]['off_flags'] + (4 if side < 0 else 0)
        d[17] = self.err.data_ptr()
        d[18] = int(float(os.environ.get('PCS_SLAB_TIMEOUT_S', '20')) * 1000)     # device-side wait timeout
        d[19] = int(os.environ.get('PCS_SLAB_TRIM', '0'))      # shorter edge segments (rows); measured: no gain, off

    def _peer_desc(self, out_index: int):
        d, m = self._desc, self.arena.meta
        for side, (ix, iz) in ((-1, (0, 1)), (+1, (2, 3))):
            if side in self.arena.remote:
                r = self.comm.rank + side
                d[ix] = self.arena.remote[side] + m[r]['off_x'][out_index]
                d[iz] = self.arena.remote[side] + m[r]['off_z'][out_index]
        self.arena.epoch += 1
        d[16] = self.arena.epoch
        return d

    # -- state ---------------------------------------------------------------------------------------------------------
    def init_state(self, x0, z0):
        """Load this rank's rows of ``x0`` / ``z0`` (None = zeros) into buffer 0, halos included."""
        from .._array import to_device
        lay, comm = self.lay, self.comm
        comm.barrier()                                   # every rank has finished its previous solve
        x, z = self.xbuf[0], self.zbuf[0]
        x.zero_()
        z.zero_()
        if x0 is not None:
            own = to_device(x0, self.dtype).view(lay.n_own, lay.W)
            ext = gather_halo(own, lay.halo_up, lay.halo_dn, comm.rank, comm.world, comm.group)
            x.view(lay.H_loc, lay.W).copy_(ext)
        if z0 is not None:
            zz = to_device(z0, self.dtype).view(2, lay.n_own, lay.W)
            for p in (0, 1):
                ext = gather_halo(zz[p], lay.halo_up, lay.halo_dn, comm.rank, comm.world, comm.group)
                z.view(2, lay.H_loc, lay.W)[p].copy_(ext)
        comm.barrier()
        return x, z

    def extract(self, x, z):
        lay = self.lay
        xo = lay.own(x).reshape(-1).clone()
        zv = z.view(2, lay.H_loc, lay.W)[:, lay.row_begin:lay.row_end]
        return xo, zv.reshape(-1).clone()

    def step(self, x, z, drow):
        s, lay = self.solver, self.lay
        if not getattr(self, '_b_ready', True):
            self.begin_solve()
        cur = 0 if x.data_ptr() == self.xbuf[0].data_ptr() else 1
        xo, zo = self.xbuf[1 - cur], self.zbuf[1 - cur]
        d = self._peer_desc(1 - cur)
        crow = self.crow.ctypes.data_as(C.c_void_p) if self.crow is not None else None
        ccol = self.ccol.ctypes.data_as(C.c_void_p) if self.ccol is not None else None
        K = self._K
        from .. import _lib
        _lib.call('pcs_pds_tv_step_slab' + self.sfx, xo.data_ptr(), zo.data_ptr(), x.data_ptr(), z.data_ptr(),
                  K.ptr(self.b), lay.H_loc, lay.W, lay.row_begin, lay.row_end, lay.grow0, lay.Hg, self.data_mode,
                  self.ntaps, crow, ccol, K.ptr(self.edge), self.nl, self.nr, self.nt, self.nb, self.g_mode, self.g_a,
                  self.g_b, self.h_mode, self.lam, float(s.tau), float(s.sigma), float(s.rho), self.steps[0],
                  self.steps[1], 0, 0, 1, drow.data_ptr(), self.ws.data_ptr(), K.stream(),
                  d.ctypes.data_as(C.c_void_p))
        if self.comm.lockstep and self.comm.world > 1:
            self.comm.barrier()                      # device synchronisation + barrier: see SlabComm.__init__
        return xo, zo

    def check(self):
        """Raise if a flag wait timed out in any step since the last check (a neighbour died or fell behind)."""
        if int(self.err[0].item()) != 0:
            self.err.zero_()
            raise RuntimeError('slab solve: a wait for a neighbour GPU timed out; the iterates are invalid')

    def close(self):
        """Release the peer arena (collective).  Optional: arenas are pooled and reused by later solves."""
        release_arenas()
