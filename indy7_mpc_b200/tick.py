"""Device-resident closed-loop MPC tick of a batch of sampled problems with every buffer bound ONCE.

The generic `Engine` calls accept any mix of numpy / torch / host / device arguments and allocate their outputs; a control
loop that runs thousands of ticks on the same buffers pays that flexibility on the host every tick (~175 us for the
four calls of a tick, against ~310 us of GPU work at 512 problems — little head-room when eight ranks share a host).
`DeviceTick` preallocates the buffers, caches their addresses and issues exactly one C-ABI call per stage:

    goal(t)    goal window of every lane out of one reference trajectory      i7_goal_window   (gato_controller.py:210-218)
    solve()    sqp() of every lane, warm-started                              i7_solve         (src/osqp_sqp.py:76-93)
    select()   per-lane costs + best trajectory: local selection, peer-memory exchange, or both fused into solve()
                                                                              i7_pack_best | i7_p2p_exchange | i7_solve_select
    shift()    warm start and x_0 of the next tick                            i7_shift         (src/osqp_mpc.py:65-70)

Results stay on the device: `out` (B, 18N-6), `stats` (B, 12), `best` ([best_cost, global index, XU_best]), `packed`
(local selection: [costs | best...]).  Everything runs on torch's current stream of the engine's device.
"""
import torch

from . import _lib


class DeviceTick:
    def __init__(self, engine, traj, offsets, xu0, x0, mode="local", index_offset=0, wrap=True):
        """mode: 'local' (i7_pack_best), 'p2p' (i7_p2p_exchange; needs connected peers) or 'fused' (i7_solve_select)."""
        if mode not in ("local", "p2p", "fused"):
            raise ValueError("mode must be 'local', 'p2p' or 'fused'")
        e = self.engine = engine
        dev = torch.device("cuda", e.device)
        tdt = e._torch_dtype()
        B, N, nz = int(xu0.shape[0]), e.N, e.nz
        if B > e.batch:
            raise ValueError(f"batch {B} exceeds the engine's maximum batch {e.batch}")
        self.B, self.mode, self.index_offset, self.wrap = B, mode, int(index_offset), int(bool(wrap))
        as_dev = lambda a, dt: torch.as_tensor(a).to(device=dev, dtype=dt).contiguous()
        self.traj = as_dev(traj, tdt)
        self.traj_len = int(self.traj.shape[0])
        self.traj_stride = self.traj.numel() // self.traj_len
        self.offsets = as_dev(offsets, torch.int32)
        self.xu = as_dev(xu0, tdt).reshape(B, nz).clone()
        self.x0 = as_dev(x0, tdt).reshape(B, 12).clone()
        self.goals = torch.empty(B, 3 * N, dtype=tdt, device=dev)
        self.out = torch.empty(B, nz, dtype=tdt, device=dev)
        self.stats = torch.empty(B, _lib.I7_STATS_LEN, dtype=tdt, device=dev)
        self.packed = torch.empty(B + 2 + nz, dtype=tdt, device=dev)
        self.best = self.packed[B:] if mode == "local" else torch.empty(2 + nz, dtype=tdt, device=dev)
        self._p = {k: getattr(self, k).data_ptr() for k in ("traj", "offsets", "xu", "x0", "goals", "out", "stats", "packed", "best")}
        self._dev = dev

    def _stream(self):
        return torch.cuda.current_stream(self._dev).cuda_stream

    def goal(self, tick):
        e, p = self.engine, self._p
        e._check(e.lib.i7_goal_window(e._h, self.B, p["traj"], self.traj_len, self.traj_stride, p["offsets"], int(tick), self.wrap, 3,
                                      p["goals"], self._stream()))

    def solve(self):
        e, p = self.engine, self._p
        if self.mode == "fused":
            e._check(e.lib.i7_solve_select(e._h, self.B, p["xu"], p["x0"], p["goals"], 3, p["out"], p["stats"], self.index_offset,
                                           p["best"], self._stream()))
        else:
            e._check(e.lib.i7_solve(e._h, self.B, p["xu"], p["x0"], p["goals"], 3, p["out"], p["stats"], self._stream()))

    def select(self):
        e, p = self.engine, self._p
        if self.mode == "local":
            e._check(e.lib.i7_pack_best(e._h, self.B, p["stats"], p["out"], self.index_offset, p["packed"], self._stream()))
        elif self.mode == "p2p":
            e._check(e.lib.i7_p2p_exchange(e._h, self.B, p["stats"], p["out"], self.index_offset, p["best"], self._stream()))

    def shift(self):
        e, p = self.engine, self._p
        e._check(e.lib.i7_shift(e._h, self.B, 1, 2, p["out"], None, None, p["xu"], p["x0"], self._stream()))

    def run(self, tick):
        """One whole tick; results in `out`, `stats`, `best`."""
        self.goal(tick)
        self.solve()
        self.select()
        self.shift()
