"""Multi-GPU host logic of the MPC-CBF path (one process per GPU, `torch.distributed`).

Two ways the path shards:

* scenario sharding — independent scenarios are split contiguously across ranks, constant tables
  are replicated, and there is NO data-path collective (`scenario_range`);
* agent sharding — ONE scenario's agents are split across ranks (BASELINE configs[3], 256 agents):
  each rank solves its own agents against the full state snapshot with `mpccbf_step_agents`, advances
  them along their curves like the reference driver with `mpccbf_follow_step` (an agent whose optimize
  failed keeps following its previous curve, MPCCBFFormationControl_example.cpp:161-221), and a single
  all-gather of the `[S][N/G][6]` next states per control step rebuilds the snapshot everywhere
  (`AgentShardedStepper`). The same loop without Python: `mpccbf_multi_run_sharded` (csrc/multi.cu). The IMPC
  iter-1 rows use the CURRENT neighbour states (ConnectivityIMPCCBF.cpp:174), so one exchange per
  control step suffices.

torch is plumbing here (device buffers, streams, the process group); the compute is the C-ABI
library. The collective works on any backend: NCCL over NVLink on the GPU box, gloo in the CPU
tests of the host logic.
"""
import torch
import torch.distributed as dist


def scenario_range(num_scenarios, rank, world):
    """Contiguous scenario shard of `rank`: sizes differ by at most one."""
    base, extra = divmod(num_scenarios, world)
    begin = rank * base + min(rank, extra)
    return begin, base + (1 if rank < extra else 0)


def agent_range(num_agents, rank, world):
    """Agent shard of `rank`. Equal shards are required (one all-gather, no padding)."""
    if num_agents % world != 0:
        raise ValueError(f"agent sharding needs num_agents ({num_agents}) divisible by the world size ({world})")
    cnt = num_agents // world
    return rank * cnt, cnt


def gather_states(local_next, out_full, world, group=None, scratch=None):
    """All-gather of agent states: local_next [S][N/G][6] on every rank -> out_full [S][N][6].

    One collective per control step. With S == 1 the gathered buffer already has the final
    layout; otherwise ranks' blocks are interleaved per scenario by one strided copy."""
    S, cnt, six = local_next.shape
    if world == 1:
        out_full.copy_(local_next)
        return out_full
    if S == 1:
        dist.all_gather_into_tensor(out_full.view(world * cnt, six), local_next.view(cnt, six).contiguous(), group=group)
        return out_full
    if scratch is None:
        scratch = torch.empty((world, S, cnt, six), dtype=local_next.dtype, device=local_next.device)
    dist.all_gather_into_tensor(scratch.view(world * S, cnt, six), local_next.contiguous(), group=group)
    out_full.view(S, world, cnt, six).copy_(scratch.permute(1, 0, 2, 3))
    return out_full


class AgentShardedStepper:
    """Closed-loop control steps of S scenarios whose N agents are split across the ranks."""

    def __init__(self, module, cfg, num_scenarios, rank, world, device, group=None):
        self.M, self.rank, self.world, self.group = module, rank, world, group
        self.S, self.N = num_scenarios, cfg.num_agents
        self.begin, self.cnt = agent_range(self.N, rank, world)
        self.h = module.Handle(cfg)
        # every shard runs the kernel variant the unsharded step would run: the result does not depend on `world`
        self.h.set_kernel_variant(self.h.auto_kernel_variant(self.S * self.N))
        n, it = self.h.n, cfg.impc_iter
        f64, i32 = torch.float64, torch.int32
        S, cnt = self.S, self.cnt
        self.ctrl = torch.zeros((S, cnt, n), dtype=f64, device=device)
        self.status = torch.zeros((S, cnt, it), dtype=i32, device=device)
        self.n_curves = torch.zeros((S, cnt), dtype=i32, device=device)
        self.objective = torch.zeros((S, cnt), dtype=f64, device=device)
        self.local_next = torch.zeros((S, cnt, 6), dtype=f64, device=device)
        self.scratch = torch.empty((world, S, cnt, 6), dtype=f64, device=device) if (world > 1 and S > 1) else None
        self.h.follow_reset(S, stream=torch.cuda.current_stream(device).cuda_stream if device.type == "cuda" else None)

    def step(self, states_full, targets_local, stream=None):
        """states_full [S][N][6] (same on every rank) is advanced IN PLACE by one control step.
        targets_local [S][N/G][3]. `stream`: a torch.cuda.Stream (default: the current one); the C-ABI calls, the
        copies and the collective are all ordered on it. Returns this rank's (ctrl_pts, status, n_curves)."""
        stream = stream if stream is not None else torch.cuda.current_stream(states_full.device)
        sp = stream.cuda_stream
        with torch.cuda.stream(stream):
            self.h.step_agents(self.S, self.begin, self.cnt, states_full, targets_local, self.ctrl, self.status,
                               self.n_curves, self.objective, stream=sp)
            self.h.follow_step(self.S, self.begin, self.cnt, states_full, self.ctrl, self.n_curves, stream=sp)
            if self.world > 1:
                self.local_next.copy_(states_full[:, self.begin:self.begin + self.cnt, :])
                gather_states(self.local_next, states_full, self.world, self.group, self.scratch)
        return self.ctrl, self.status, self.n_curves

    def close(self):
        self.h.close()
