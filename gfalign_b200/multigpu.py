"""`gfalign search` / `gfalign evalGFA` over the GPUs of one box: one process per GPU (torchrun), the
alignment table sharded by contiguous CSR ranges, the small count records combined with ONE NCCL
all-reduce per batch (search) / per run (evalGFA).  Same options and stdout as the single-GPU CLI
`gfalign-b200` (and therefore as the reference); rank 0 prints.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \\
      -m gfalign_b200.multigpu search -f g.gfa -g a.gaf -n nodes.tsv -s SRC -d DST [-m STEPS]
      [--min-nodes K] [--return-all-paths]
  ... -m gfalign_b200.multigpu evalGFA -f g.gfa -g a.gaf -o decorated.gfa

The heap replay that defines the row order (reference src/eval.cpp:134-189) is NOT sharded: every
rank replays it on the same summed counts (gfe_search_with_scorer), so all ranks stay in lock step.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

from . import capi, sharded, synth


def _args(argv):
    ap = argparse.ArgumentParser(prog="gfalign_b200.multigpu")
    ap.add_argument("tool", choices=["search", "evalGFA"])
    ap.add_argument("-f", "--input-sequence", dest="gfa", required=True)
    ap.add_argument("-g", "--input-alignment", dest="gaf", default=None)
    ap.add_argument("-n", "--node-file", dest="node_file", default=None)
    ap.add_argument("-s", "--source", default="")
    ap.add_argument("-d", "--destination", default="")
    ap.add_argument("-m", "--max-steps", dest="max_steps", type=int, default=100000)
    ap.add_argument("-o", "--out-format", dest="out", default=None)
    ap.add_argument("-j", "--threads", type=int, default=0)
    ap.add_argument("--min-nodes", type=int, default=0)
    ap.add_argument("--return-all-paths", action="store_true")
    ap.add_argument("--speculate", type=int, default=0)
    ap.add_argument("--shard", choices=["alignments", "batch", "auto"], default="auto",
                    help="search: shard the alignment table over the GPUs, or replicate it and deal each batch's "
                         "speculation subtrees (small tables); auto = batch up to 4 M alignments")
    return ap.parse_args(argv)


def _node_budget(tp, node_file, source, destination):
    """NodeTable + the two add() calls (reference include/nodetable.h, src/eval.cpp:125-128)."""
    ids = {nm: i for i, nm in enumerate(tp.names)}
    nodelist = []
    if node_file:
        for ln in open(node_file).read().splitlines():
            c = ln.split("\t")
            count = int(c[1]) if len(c) > 1 else 1
            if count < 1:
                continue
            if c[0] not in ids:
                sys.stderr.write(f"Error: node not in graph (pIUd: {c[0]})\n")
                sys.exit(1)
            nodelist.append((ids[c[0]], count))
    t = synth.Tangle(tp.names, tp.edge_from, tp.edge_to, source, destination, nodelist)
    return synth.node_table(t)


def _write_decorated(gfa_path, out_path, weights):
    """Same record order as csrc/host/modes.cpp: H, S, L + RC:i, then everything else."""
    lines = [l for l in open(gfa_path).read().split("\n") if l]
    with open(out_path, "w") as f:
        f.write("H\tVN:Z:1.2\n")
        f.writelines(l + "\n" for l in lines if l[0] == "S")
        k = 0
        for l in lines:
            if l.startswith("L\t"):
                f.write(f"{l}\tRC:i:{int(weights[k])}\n")
                k += 1
        for l in lines:
            if l[0] in "SLH":
                continue
            if l[0] == "P":
                c = l.split("\t")
                if len(c) >= 4:
                    c[3] = "*"
                    l = "\t".join(c)
            f.write(l + "\n")


def main(argv=None):
    a = _args(sys.argv[1:] if argv is None else argv)
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("gfalign_b200.multigpu needs CUDA devices: there is no CPU fallback")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        # NCCL prints its version banner with printf: keep our stdout clean by giving the C level
        # stderr as fd 1 and writing our own output to a private copy of the real stdout
        sys.stdout.flush()
        real = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)
        sys.stdout = real
        dist.init_process_group("nccl", device_id=device)

    tp = capi.TextPack(a.gfa, a.gaf, threads=a.threads)
    batch_shard = a.tool == "search" and (a.shard == "batch" or (a.shard == "auto" and tp.A <= 4_000_000))
    if batch_shard:
        off, st = tp.offsets, tp.steps                 # every rank keeps the whole table
    else:
        cuts = sharded.shard_bounds(tp.offsets, world)
        off, st = sharded.take_shard(tp.offsets, tp.steps, int(cuts[rank]), int(cuts[rank + 1]))
    if os.environ.get("GFE_TEST_LEGACY_RANK") == str(rank):
        # test hook: THIS rank alone runs without the pair dictionary / incremental path, as a rank whose
        # shard overflows the dictionary would; gfe_search_sharded must then use full scans on every rank
        os.environ["GFE_FORCE_LEGACY"] = "1"
    eng = capi.Engine(local)
    eng.set_alignments(off, st, tp.n_nodes)

    if a.tool == "search":
        src = max(tp.node_id(a.source), 0)          # headersToIds[...] default-inserts 0
        dst = max(tp.node_id(a.destination), 0)
        nodes, counts, node_count = _node_budget(tp, a.node_file, src, dst)
        eng.set_universe(nodes)

        def local_score(po, ps):
            return eng.score_paths(po, ps, filter=True)

        def local_replicated(po, ps):      # batch sharding without the peer kernel: every rank scores everything
            return eng.score_paths(po, ps, filter=True)

        out = sys.stdout

        def on_row(row):
            if rank == 0:
                out.write(row.columns(tp.names) + "\n")
                out.flush()

        kw = dict(max_steps=a.max_steps, min_nodes=a.min_nodes, return_all_paths=a.return_all_paths,
                  speculate=a.speculate, on_row=on_row)
        px = None
        scorer = local_score
        if world > 1:
            scorer = local_replicated if batch_shard else sharded.ShardedScorer(local_score, device=device)
            if os.environ.get("GFE_EXCHANGE", "peer") == "peer":
                try:      # counts stay on the device, summed by the peer-memory kernel (csrc/peer.cu)
                    px = sharded.PeerExchange(1 << 18, device)     # 4 slots per candidate of a 4096-wide batch + padding
                except Exception as ex:
                    sys.stderr.write(f"[multigpu] peer exchange unavailable ({ex!r}); using NCCL\n")
        if px is not None:
            # the whole loop in C: gfe_search_sharded -> gfe_score_paths_incremental_sharded per batch
            if batch_shard:
                px.c_exchange().set_batch_sharding(True)
            rows, stats = eng.search_sharded(px.c_exchange(), tp.adj_offsets, tp.adj_target, tp.adj_or0,
                                             tp.adj_or1, src, dst, nodes, counts, node_count, **kw)
        else:
            rows, stats = capi.search_with_scorer(scorer, tp.adj_offsets, tp.adj_target, tp.adj_or0,
                                                  tp.adj_or1, src, dst, nodes, counts, node_count, **kw)
        if rank == 0 and os.environ.get("GFE_DEBUG_TIMING"):
            sys.stderr.write(f"[multigpu] {'candidate-batch' if batch_shard else 'alignment'} sharding, "
                             f"exchange {'peer kernel (C loop)' if px is not None else getattr(scorer, '__name__', type(scorer).__name__)}, "
                             f"{stats['kernel_batches']} batches, scoring {stats['seconds_scoring']:.3f} s\n")
        if rank == 0 and stats["extractions"] >= a.max_steps:
            out.write(f"Reached maximum number of steps ({stats['extractions']})\n")
    else:
        if rank == 0 and a.gaf:
            # the alignment summary evalGFA prints first is host-only code of the CLI binary
            import subprocess
            cli = os.path.join(os.path.dirname(os.path.abspath(__file__)), "bin", "gfalign-b200")
            sys.stdout.write(subprocess.run([cli, "evalGFA", "-f", a.gfa, "-g", a.gaf], capture_output=True,
                                            text=True, check=True).stdout)
            sys.stdout.flush()
        E = int(tp.edge_from.size)
        d = torch.zeros(max(2 * E, 1), dtype=torch.int32, device=device)
        px = None
        if world > 1 and E and os.environ.get("GFE_EXCHANGE", "peer") == "peer":
            try:      # the 2E-slot record summed by the same peer-memory kernel as the search records
                px = sharded.PeerExchange(2 * E, device)
            except Exception as ex:
                sys.stderr.write(f"[multigpu] peer exchange unavailable ({ex!r}); using NCCL\n")
        if px is not None:
            stream = torch.cuda.current_stream(device).cuda_stream
            eng.edge_traversals_device(tp.edge_from, tp.edge_to, px.half_ptr(), stream)
            px.sum_into(d.data_ptr(), 2 * E, stream)
            torch.cuda.synchronize()
        else:
            eng.edge_traversals_device(tp.edge_from, tp.edge_to, d.data_ptr())
            torch.cuda.synchronize()
            if world > 1:
                dist.all_reduce(d)                   # integer sum over the alignment shards
        if rank == 0 and a.out:
            w = capi.Engine.edge_weights_from_counts(tp.edge_from, tp.edge_to,
                                                     d[:2 * E].cpu().numpy().astype(np.uint32))
            _write_decorated(a.gfa, a.out, w)
    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
