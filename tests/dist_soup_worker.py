"""torchrun worker for the sharded-soup tests: every rank owns a contiguous block of particles of a PopulationSoup,
runs a few generations (one all-gather of the record arena per generation) and rank 0 saves the gathered result.

  --backend gloo : CPU; PopulationSoup's real host logic (sharding, all-gather, uid offsets, count all-reduce) with
                   the ORACLE standing in for the kernels (test-only backend defined here)
  --backend nccl : one rank per GPU, the real engine
"""
import argparse
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "self-replicating-neural-networks_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


class OracleBackend:
    """CPU stand-in for the engine (tests only): same calls, numpy views of CPU tensors, oracle arithmetic."""
    device = torch.device("cpu")

    def alloc(self, shape, dtype):
        return torch.zeros(shape, dtype=dtype)

    @staticmethod
    def _odesc(desc):
        import oracle
        return oracle.desc(desc.kind, desc.width, desc.depth, desc.aggregates)

    def init(self, desc, w, seed, slot0, generation):
        import oracle
        oracle.init(self._odesc(desc), w.numpy(), seed, slot0=slot0, generation=generation)

    def soup_sync(self, desc, sp, snap, n_total, w_out, slot0, n, seed, gen, scratch, action, dead):
        import oracle
        sp = {k: v for k, v in sp.items() if k != "engine_path"}
        a, _, _ = oracle.soup_sync(self._odesc(desc), oracle.soup_params(**sp), snap.numpy(), n_total, w_out.numpy(),
                                   slot0, n, seed, gen)
        action.copy_(torch.from_numpy(a))
        dead.copy_(torch.from_numpy((a >= 4).astype(np.uint8)))

    def classify(self, desc, w, eps, with_fix_sec):
        import oracle
        return torch.from_numpy(oracle.classify(self._odesc(desc), w.numpy(), eps, with_fix_sec)[1])

    def assign_uids(self, dead, uid, base, tmp, n_dead):
        d = dead.numpy().astype(bool)
        uid.numpy()[d] = base + np.arange(int(d.sum()))
        n_dead[0] = int(d.sum())


def run_together(args, rank, world):
    """A mixed soup: the three kinds advance in lock step, one cross-GPU barrier per generation."""
    from soup import PopulationSoup, evolve_together
    kw = dict(attacking_rate=0.3, learn_from_rate=0.2, train=1, learn_from_severity=2, remove_divergent=True,
              remove_zero=True, epsilon=1e-4)
    soups = [PopulationSoup(k, args.size, aggregates=4 if k == 1 else 0, seed=31, exchange=args.exchange, **kw)
             for k in (0, 1, 2)]
    for ps in soups:
        ps.track_uids = True
        ps.seed_particles()
    evolve_together(soups, args.gens)
    out = dict(world=world, exchange=soups[0].exchange)
    for k, ps in enumerate(soups):
        counts = ps.count()
        full = torch.zeros((ps.size, ps.desc.n_weights), dtype=torch.float32, device=ps.w.device)
        uids = torch.zeros(ps.size, dtype=torch.int64, device=ps.w.device)
        dist.all_gather_into_tensor(full, ps.w.contiguous())
        dist.all_gather_into_tensor(uids, ps.uid)
        out.update({"w%d" % k: full.cpu().numpy(), "uid%d" % k: uids.cpu().numpy(),
                    "counts%d" % k: np.array(list(counts.values())), "next_uid%d" % k: ps.next_uid, "time%d" % k: ps.time})
    if rank == 0:
        np.savez(args.out, **out)
    dist.barrier()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", required=True)
    ap.add_argument("--out", required=True)
    ap.add_argument("--kind", type=int, default=0)
    ap.add_argument("--size", type=int, default=4096)
    ap.add_argument("--gens", type=int, default=4)
    ap.add_argument("--exchange", default="auto")
    ap.add_argument("--together", type=int, default=0, help="1: all three kinds in lock step (soup.evolve_together)")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    if args.backend == "nccl":
        torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
        dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    else:
        dist.init_process_group("gloo")
    from soup import PopulationSoup
    if args.together:
        run_together(args, rank, world)
        return
    ps = PopulationSoup(args.kind, args.size, aggregates=4 if args.kind == 1 else 0, seed=31,
                        backend=OracleBackend() if args.backend == "gloo" else None, exchange=args.exchange,
                        attacking_rate=0.3,
                        learn_from_rate=0.2, train=1, learn_from_severity=2, remove_divergent=True, remove_zero=True,
                        epsilon=1e-4)
    ps.track_uids = True
    ps.seed_particles()
    ps.evolve(args.gens)
    counts = ps.count()
    full = torch.zeros((ps.size, ps.desc.n_weights), dtype=torch.float32, device=ps.w.device)
    uids = torch.zeros(ps.size, dtype=torch.int64, device=ps.w.device)
    dist.all_gather_into_tensor(full, ps.w.contiguous())
    dist.all_gather_into_tensor(uids, ps.uid)
    comm_ok = -1
    if args.backend == "nccl":
        # the C ABI's own communicator (srnn_comm_*): same all-gather without torch.distributed on the data path
        import srnn_engine as eng
        box = [eng.Comm.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        comm = eng.Comm(box[0], rank, world)
        again = torch.zeros_like(full)
        comm.allgather(ps.w.contiguous(), again)
        torch.cuda.synchronize()
        comm_ok = int(torch.equal(again.view(torch.int32), full.view(torch.int32)))
        comm.destroy()
    if rank == 0:
        np.savez(args.out, w=full.cpu().numpy(), uid=uids.cpu().numpy(), counts=np.array(list(counts.values())),
                 next_uid=ps.next_uid, world=world, exchange=ps.exchange, comm_ok=comm_ok)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
