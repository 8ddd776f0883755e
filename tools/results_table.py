"""profiles/r02_results.md from the JSON lines of the end-of-round runs (tools/final_check.sh on one GPU, the 2- and 8-GPU
`torchrun` lines): python tools/results_table.py > profiles/r02_results.md"""
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")


def load(name):
    for line in open(os.path.join(OUT, name)):
        if line.startswith("{"):
            return json.loads(line)
    raise SystemExit("no JSON line in " + name)


def M(v):
    return "%.1f M" % (v / 1e6)


def head(b, r, n2, n8, r8, c):
    o = ["# Round 2 measured results (B200, gpurun), end of round\n",
         "All lines are `bench.py` outputs of the committed build (`tools/final_check.sh` on one GPU; the 2- and 8-GPU lines "
         "under `torchrun`; JSON kept under `gpurun_out/`, scratch; this file: `tools/results_table.py`).  C2 is the contract "
         "line, the rest is supplementary.  In brackets: the same line at the START of round 2's kernel work (= the state the "
         "round-1 verdict saw plus this round's parity / boundary work; round 1 itself: `profiles/r01_results.md`).\n",
         "| configuration | agent-steps/s | ms per iteration | notes |\n|---|---|---|---|"]
    o.append("| C2: simple_spread, 3 agents, 65,536 envs, shared MLP, 1 GPU (`python bench.py --steps 20 --warmup 5`) | %s (138.5 M) "
             "| %.2f (35.50) | e2e through host arrays %s (82.2 M; round 1: 66 – 69 M): H2D 20 MB + D2H 501 MB per iteration |"
             % (M(b["value"]), b["ms_per_step"], M(b["e2e"]["value"])))
    o.append("| the UNMODIFIED reference on the same box (`--impl reference`: SubprocVecEnv x 128 + MPERunner + R_MAPPO, torch CPU, "
             "%d cores) | %.1f k | %.0f | `cpu_baseline.kind = \"reference\"` (%.1f k inside the C2 line); the CPU port "
             "(`cpu_baseline_port`): %.0f k |" % (r["cpu_baseline"]["cores"], r["value"] / 1e3, r["ms_per_step"],
                                                  b["cpu_baseline"]["value"] / 1e3, b["cpu_baseline_port"]["value"] / 1e3))
    c1 = b["c1"]
    o.append("| C1 (the reference arm's own configuration: 128 envs) on 1 GPU, `c1` key | %.1f M eager / %.1f M CUDA graph | "
             "%.2f / %.2f (3.41 / 2.15) | same configuration on both arms: x %.0f (graph) |"
             % (c1["eager"]["value"] / 1e6, c1["cuda_graph"]["value"] / 1e6, c1["eager"]["ms_per_step"],
                c1["cuda_graph"]["ms_per_step"], c1["cuda_graph"]["value"] / r["value"]))
    o.append("| C2 on 2 GPUs | %s (276.5 M) | %.2f (35.56) | e2e %s (153.1 M); one float64 all-reduce per update |"
             % (M(n2["value"]), n2["ms_per_step"], M(n2["e2e"]["value"])))
    o.append("| C2 on 8 GPUs (`torchrun --nproc-per-node 8 bench.py --gpus 8 --steps 20 --warmup 5 --c5`) | %s (1,105.3 M) | "
             "%.2f (35.58) | %.3f of 8 x 1 GPU; e2e %s (466.9 M); reference arm on that box (%d cores): %.1f k |"
             % (M(n8["value"]), n8["ms_per_step"], n8["value"] / (8 * b["value"]), M(n8["e2e"]["value"]),
                r8["cpu_baseline"]["cores"], r8["value"] / 1e3))
    c5 = n8["c5"]
    o.append("| C5 on 8 GPUs (`c5` key of the same line): simple_spread, 10 agents, 16,384 envs per GPU = %s concurrent agents | "
             "%s (932.8 M; round 1: 729 M) | %.2f (35.13) | north_star target (>= 1 M concurrent agents) |"
             % (format(c5["concurrent_agents"], ","), M(c5["value"]), c5["ms_per_step"]))
    o.append("| C5 with 65,536 envs x 10 agents on 1 GPU (`--config c5`) | %s (133.5 M) | %.1f (122.7) | e2e %s |"
             % (M(c["c5"]["value"]), c["c5"]["ms_per_step"], M(c["c5"]["e2e"]["value"])))
    o.append("| C4: simple_reference, shared GRU policy, 32,768 envs, 1 GPU (`--config c4`) | %s (9.7 M) | %.1f (169.5) | one "
             "weight-gradient pass per GRU layer instead of two (167.3 ms before it); `profiles/r02_c4_launches.md`, "
             "`profiles/r02_c4_kernels.md`, DESIGN 4.3 |" % (M(c["c4"]["value"]), c["c4"]["ms_per_step"]))
    o.append("| C3: speaker_listener, separated MLP policies, 32,768 envs (`--config c3`) | %s (48.4 M) | %.1f (33.8) | with GRU "
             "policies (`c3r`): %s, %.1f ms (8.6 M, 189.9) |"
             % (M(c["c3"]["value"]), c["c3"]["ms_per_step"], M(c["c3r"]["value"]), c["c3r"]["ms_per_step"]))
    return o


def tables(b, n2, n8):
    o = ["\n## Per-kernel roofline, C2, 1 GPU (`roofline.kernels` of the contract line)\n",
         "| kernel | launches / iteration | ms / launch | ms / iteration | algorithmic work / launch | achieved | fraction of "
         "measured peak | DRAM traffic / launch (ncu) |\n|---|---|---|---|---|---|---|---|"]
    tot = 0.0
    for k in b["roofline"]["kernels"]:
        tr = k.get("traffic")
        o.append("| %s | %g | %.4f | %.3f | %s | %.1f %s | %.4f (%s) | %s |"
                 % (k["kernel"], k["launches_per_step"], k["ms_per_launch"], k["ms_per_step"], k["work_formula"], k["achieved"],
                    k["unit"], k["frac"], "bf16 sustained" if k["bound"] == "tensor" else "HBM copy",
                    ("%.1f MB" % (tr / 1e6)) if tr else "-"))
        tot += k["ms_per_step"]
    o.append("\nThe entries add up to %.1f %% of the iteration.  Peaks: `MEASURED_PEAKS.json` (HBM 6464.6 GB/s, bf16 sustained "
             "1400.8 TFLOP/s).  Start of the round: actor update 2.208 ms, critic update 0.805 ms, actor forward 77 us, critic "
             "forward 34 us (DESIGN.md 4.1 has the change-by-change table; ncu before / after: `profiles/r02_mlp_kernels.md` / "
             "`profiles/r02_mlp_kernels_b.md`; launch lists `profiles/r02_c2_launches_a.md` / `_b.md`).\n"
             % (100 * tot / b["ms_per_step"]))
    o.append("## C5 on 8 GPUs, per kernel (`c5.kernels`)\n\n| kernel | ms / launch | ms / iteration | fraction of measured peak |"
             "\n|---|---|---|---|")
    for k in n8["c5"]["kernels"]:
        o.append("| %s | %.4f | %.3f | %.4f |" % (k["key"], k["ms_per_launch"], k["ms_per_step"], k["frac"]))
    o.append("\n## Multi-rank parity (`dp_check`)\n\n| ranks | replicas identical | gradient vs single GPU on the union (max / "
             "max|g|) | scalars | parameters (abs) | identical after the timed region |\n|---|---|---|---|---|---|")
    for d in (n2, n8):
        q = d["dp_check"]
        o.append("| %d | %s | %.1e | %.1e | %.1e | %s |"
                 % (q["ranks"], q["replicas_identical"], q["grad_vs_single_gpu_max_rel"], q["info_vs_single_gpu_max_rel"],
                    q["params_vs_single_gpu_max_abs"], q["replicas_identical_after_timed_region"]))
    o.append("\n## Tests on the B200\n\n`python -m pytest tests -q -m gpu`: 185 passed, 1 skipped on one GPU (the 2-rank test needs "
             "2 GPUs and passes under `gpurun --gpus 2`); `tests/test_zz_no_alloc.py` passes alone and last.  Large-batch gradient "
             "distances with the end-of-round kernels (`tests/test_gpu_large_batch.py`, worst |dg| / max|g| per case): actor "
             "8.3e-6, critic (group 3) 1.1e-6, ReLU group 1 5.5e-6 / 2.5e-6, wide critic packed 1.3e-5, unpacked 3.4e-6, packed "
             "ReLU 1.7e-5, GRU Tanh 1.8e-6 / 1.9e-6, GRU ReLU 6.3e-5 / 7.8e-5 (tolerance 1e-4) -- the uniformly scaled operand "
             "pairs sit where the 2^12-scaled pairs sat.  `tests/test_gpu_bulk_paths.py`: the bulk-copy and plain-load paths of "
             "both update kernels give bit-identical sums.")
    return o


def main():
    b, r = load("final_bench.json"), load("final_ref.json")
    n2, n8, r8 = load("final_n2.json"), load("final_n8.json"), load("final_ref_n8.json")
    c = {k: load("final_%s.json" % k) for k in ("c3", "c3r", "c4", "c5")}
    print("\n".join(head(b, r, n2, n8, r8, c) + tables(b, n2, n8)))


if __name__ == "__main__":
    main()
