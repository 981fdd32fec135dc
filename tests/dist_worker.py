"""Worker for the multi-process tests: run under torchrun.
   gloo mode (CPU): exercises the host-side plan (splitters, count matrix, receive layout)
                    with numpy standing in for the device kernels.
   nccl mode (GPU): runs DistributedSorter end to end and checks it against the oracle."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as o  # noqa: E402


def shard(rank, n, kind):
    x = o.pcg_stream((rank + 1) * n)[rank * n:]
    if kind == "consthigh":
        x = (np.uint32(0xAB000000) | (x & np.uint32(0x00FFFFFF))).astype(np.uint32)
    elif kind == "zipf":
        x = (np.floor(2.0 ** (32.0 * x.astype(np.float64) / 2.0 ** 32)) - 1).astype(np.uint32)
    elif kind == "fewvalues":
        x = (x % np.uint32(5)).astype(np.uint32) * np.uint32(1000)
    elif kind == "allequal":
        x = np.full(x.size, 0x12345678, np.uint32)
    return x


def main():
    mode, n = sys.argv[1], int(sys.argv[2])
    plan = importlib.import_module("gpu-radix-sort_b200.distrib_plan")
    if mode == "gloo":
        dist.init_process_group("gloo")
        rank, world = dist.get_rank(), dist.get_world_size()
        for kind in ("uniform", "consthigh", "zipf", "fewvalues", "allequal"):
            keys = shard(rank, n, kind)
            sample = torch.from_numpy(keys[:: max(1, n // 256)][:256].astype(np.int64))
            gathered = [torch.empty_like(sample) for _ in range(world)]
            dist.all_gather(gathered, sample)
            spl, thr = plan.choose_splitters(np.sort(torch.cat(gathered).numpy().astype(np.uint32)), world, with_ties=True)
            dest = plan.destination(keys, spl, thr)          # ties cut by position in the shard
            cnt = torch.from_numpy(np.bincount(dest, minlength=world).astype(np.int64))
            rows = [torch.empty_like(cnt) for _ in range(world)]
            dist.all_gather(rows, cnt)
            cm = torch.stack(rows).numpy()
            offsets, totals = plan.recv_layout(cm)
            assert totals.sum() == n * world and (offsets[0] == 0).all()
            assert (offsets[-1] + cm[-1] == totals).all()
            # exchange with all_to_all, then local sort: rank r must hold the r-th value range
            order = np.argsort(dest, kind="stable")
            send = [keys[order][dest[order] == r] for r in range(world)]
            everyone = [None] * world                      # gloo has no all-to-all: gather and pick our column
            dist.all_gather_object(everyone, send)
            recv = [everyone[s][rank] for s in range(world)]
            assert [len(x) for x in recv] == cm[:, rank].tolist()
            mine = np.sort(np.concatenate(recv).astype(np.uint32))
            allkeys = np.sort(np.concatenate([shard(r, n, kind) for r in range(world)]))
            start = int(totals[:rank].sum())
            assert np.array_equal(mine, allkeys[start:start + mine.size]), kind
            if kind in ("fewvalues", "allequal"):
                assert plan.imbalance(totals) < 1.1, (kind, totals)
        # step-wise mode: counts -> all-gather -> bases -> every key to its global position's owner
        for kind, lens in (("uniform", [n] * world), ("zipf", [n + 17 * r for r in range(world)]), ("allequal", [n] * world)):
            mine = shard(rank, lens[rank], kind)
            whole = np.concatenate([shard(r, lens[r], kind) for r in range(world)])
            starts = np.concatenate([[0], np.cumsum(lens)])
            for step in range(4):
                digit = (mine >> np.uint32(8 * step)) & np.uint32(0xFF)
                cnt = torch.from_numpy(np.bincount(digit, minlength=256).astype(np.int64))
                rows = [torch.empty_like(cnt) for _ in range(world)]
                dist.all_gather(rows, cnt)
                bases = plan.step_bases(torch.stack(rows).numpy(), rank)
                pos = plan.step_positions(mine, 8 * step, 8, bases)
                owner, off = plan.slice_of(pos, starts)
                sent = [(off[owner == r], mine[owner == r]) for r in range(world)]
                everyone = [None] * world
                dist.all_gather_object(everyone, sent)
                nxt = np.zeros(lens[rank], dtype=np.uint32)
                filled = 0
                for src in range(world):
                    o, v = everyone[src][rank]
                    nxt[o] = v
                    filled += len(o)
                assert filled == lens[rank]
                whole = whole[np.argsort((whole >> np.uint32(8 * step)) & np.uint32(0xFF), kind="stable")]
                assert np.array_equal(nxt, whole[starts[rank]:starts[rank + 1]]), (kind, "step", step + 1)
                mine = nxt
            assert np.array_equal(mine, np.sort(whole)[starts[rank]:starts[rank + 1]])
        print("rank %d gloo plan ok" % rank)
    else:
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        import datetime
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=120))
        rank, world = dist.get_rank(), dist.get_world_size()
        distrib = importlib.import_module("gpu-radix-sort_b200.distrib")
        for exch in ("p2p", "nccl"):
            sorter = distrib.DistributedSorter(n, mode=exch, slack=2.5)
            for kind in ("uniform", "consthigh", "zipf", "fewvalues", "allequal"):
                keys = shard(rank, n, kind)
                d = torch.from_numpy(keys.view(np.int32)).cuda()
                for rep in range(2):
                    out, cnt = sorter.sort(d)
                stats = sorter.verify(d, out, cnt)
                allkeys = np.sort(np.concatenate([shard(r, n, kind) for r in range(world)]))
                start = int(sum(stats["keys_per_rank"][:rank]))
                got = out.cpu().numpy().view(np.uint32)
                assert np.array_equal(got, allkeys[start:start + cnt]), (exch, kind)
                if kind in ("fewvalues", "allequal"):
                    assert stats["imbalance"] < 1.1, (kind, stats)
                if rank == 0:
                    print(exch, kind, "ok; keys/rank", stats["keys_per_rank"], "imbalance %.3f" % stats["imbalance"])
            if exch == "p2p":
                # the reference's step-wise protocol: intermediate state after step k == the whole array
                # stably sorted by its low 8k bits, cut back into the ranks' slices
                for kind in ("uniform", "zipf", "allequal"):
                    keys = shard(rank, n, kind)
                    d = torch.from_numpy(keys.view(np.int32)).cuda()
                    allkeys = np.concatenate([shard(r, n, kind) for r in range(world)])
                    seen = []
                    out, cnt = sorter.sort_stepwise(d, on_step=lambda k, v: seen.append(v.cpu().numpy().view(np.uint32).copy()))
                    assert cnt == n and len(seen) == 4
                    state = allkeys
                    for k in range(4):
                        state = state[np.argsort((state >> np.uint32(8 * k)) & np.uint32(0xFF), kind="stable")]
                        assert np.array_equal(seen[k], state[rank * n:(rank + 1) * n]), (kind, "step", k + 1)
                    assert np.array_equal(out.cpu().numpy().view(np.uint32), np.sort(allkeys)[rank * n:(rank + 1) * n]), kind
                    sorter.verify(d, out, cnt)
                    if rank == 0:
                        print("stepwise", kind, "ok")
            sorter.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
