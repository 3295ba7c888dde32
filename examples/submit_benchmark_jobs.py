#!/usr/bin/env python
"""Scaling sweep launcher: runs ``examples/benchmark.py`` at 1..N ranks.

Counterpart of the reference's SLURM launcher
(``tests/smf_example/submit_benchmark_jobs.py``: ten sbatch files, 1..10 MPI
tasks).  The target here is one NVSwitch box, so by default the sweep is run
locally through ``torchrun`` (1, 2, 4, 8 GPUs); ``--slurm`` writes one sbatch
script per size instead (``srun torchrun ...`` on a single node).
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))  # _bootstrap: run from the source tree
import argparse
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent

parser = argparse.ArgumentParser(description=__doc__.split("\n")[0])
parser.add_argument("--num-halos", type=int, default=100_000)
parser.add_argument("--num-steps", type=int, default=100)
parser.add_argument("--sizes", type=int, nargs="+", default=[1, 2, 4, 8])
parser.add_argument("--save", default="benchmark_results.txt")
parser.add_argument("--slurm", action="store_true")
parser.add_argument("--partition", default="gpu")
parser.add_argument("--account", default=None)


def command(n, args, port):
    return [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
            f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
            "--master-port", str(port), str(HERE / "benchmark.py"),
            "--num-halos", str(args.num_halos), "--num-steps",
            str(args.num_steps), "--save", args.save]


def main():
    args = parser.parse_args()
    for i, n in enumerate(args.sizes):
        cmd = command(n, args, 29700 + i)
        if args.slurm:
            lines = ["#!/bin/bash", f"#SBATCH --job-name=mgb200_bench_{n}",
                     f"#SBATCH --partition={args.partition}",
                     "#SBATCH --nodes=1", f"#SBATCH --gpus-per-node={n}",
                     f"#SBATCH --ntasks-per-node=1",
                     f"#SBATCH --output=bench_{n}.out"]
            if args.account:
                lines.append(f"#SBATCH --account={args.account}")
            lines.append("srun " + " ".join(cmd))
            path = Path(f"bench_job_{n}.sbatch")
            path.write_text("\n".join(lines) + "\n")
            print("wrote", path, "-> submit with: sbatch", path)
        else:
            print(">>", " ".join(cmd), flush=True)
            subprocess.run(cmd, check=True)


if __name__ == "__main__":
    main()
