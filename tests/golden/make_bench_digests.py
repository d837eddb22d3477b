"""Golden state digests for bench.py's `parity_check` (test infrastructure; run where the oracle builds):

    python tests/golden/make_bench_digests.py [C2:25 C2:105 ...]

For every `workload:frames` it runs the CPU checker (oracle/librestate.so, pinned to the reference CPU engines by
tests/test_oracle_vs_ref.py) over the first `frames` frames of the workload's synthetic sequence with bench.py's capacities
and per-frame schedule (process_frame, then SlideWindow + Decay where the workload prunes), and stores the digest of the
final state (denseslam-global-consistency-h_b200/digest.py) in tests/golden/bench_digests.json.  bench.py never runs the
oracle for this check: it compares its own digest with the committed one."""
import importlib
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
import harness as hz  # noqa: E402
import bench  # noqa: E402

digest = importlib.import_module(hz.PKG + ".digest")
OUT = os.path.join(HERE, "bench_digests.json")
DEFAULT = ["C2:25", "C2:105", "C1:25", "C1:105", "C3:25", "C3:105"]


def oracle_digest(workload, n_frames):
    spec = bench.WORKLOADS[workload]
    cfg, frames = hz.synth.make_sequence(workload, n_frames, seed=0)
    ora = hz.OracleBackend(cfg, spec["caps"])
    ora.set_weights(cfg.depth_weighting, 10, 100)
    prune = spec["prune"]
    for f in frames:
        ora.set_frame(f)
        assert ora.process_frame() == 0
        if prune:
            ora.slide_window(prune["max_age"])
            ora.decay(prune["max_weight"], prune["min_age"])
    s = ora.state(voxels=False)
    vox = ora.scene.voxels
    d = digest.state_digest(s["hash"], s["last_free"], s["visible"], lambda lo, hi: vox[lo * 512:hi * 512],
                            ora.scene.c.num_blocks, ora.images())
    d["no_visible"] = int(len(s["visible"]))
    ora.close()
    return d


def main():
    want = sys.argv[1:] or DEFAULT
    table = json.load(open(OUT)) if os.path.exists(OUT) else {}
    for item in want:
        workload, n = item.split(":")
        t0 = time.time()
        table[item] = oracle_digest(workload, int(n))
        print("%s: %d blocks, %d visible, %.1f s" % (item, table[item]["blocks_allocated"], table[item]["no_visible"], time.time() - t0), flush=True)
        json.dump(table, open(OUT, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
