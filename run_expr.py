"""Experiment sweep — same role, CSV schema and regex as the reference's run_expr.py
(block size x iteration count, matrix = 8 * block, one CSV column per method scraped from
the "average time" lines, run_expr.py:6-43), retargeted from `bsub` on Sunway to the local
stencil_main binary with the current flag names (-s size, -m methods, -R repeats).
Extra columns: GStencil/s per method.  Run on a GPU box (or under gpurun)."""
import argparse
import csv
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(ROOT, "stencil_b200", "host", "build", "bin", "stencil_main")

block_sizes = [8, 16, 32, 48, 50, 64, 72, 88, 100, 120]
iterations = [1, 10, 100, 1000, 5000, 10000, 50000, 100000]
pattern = r'The average time taken by (.*) method is (.*)ms for \d+ iterations\.'
rate_pattern = r'^\s+(\S+) sweep .* ms \(best\), (\S+) GStencil/s'


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--methods", nargs="+", default=["CUDA"])
    ap.add_argument("--repeat", type=int, default=3)
    ap.add_argument("--out", default="output.csv")
    ap.add_argument("--blocks", nargs="+", type=int, default=block_sizes)
    ap.add_argument("--iters", nargs="+", type=int, default=iterations)
    ap.add_argument("--size", type=int, default=0,
                    help="fixed matrix size (default: 8 * block, as in the reference's sweep, run_expr.py:14)")
    ap.add_argument("--block-as-tile", action="store_true",
                    help="CUDA reads -b as its tile preference (rows per warp), so the block column of the CSV "
                         "is a real tuning axis on this backend too; combine with --size to sweep it at a fixed N")
    ap.add_argument("--extra", nargs=argparse.REMAINDER, default=[], help="further stencil_main flags")
    args = ap.parse_args()
    outputs = []
    for block_size in args.blocks:
        for iteration in args.iters:
            matrix_size = args.size or block_size * 8
            command = [BIN, '-s', str(matrix_size), '-b', str(block_size), '-i', str(iteration),
                       '-R', str(args.repeat), '-m', *args.methods]
            if args.block_as_tile:
                command.append('--block-as-tile')
            command += args.extra
            print(f'----------- block size: {block_size}, iteration: {iteration} -----------')
            result = subprocess.run(command, capture_output=True, text=True)
            output = result.stdout.strip()
            print(output)
            result_dict = {'Block Size': block_size, 'Iteration': iteration}
            for match in re.finditer(pattern, output):
                method, time = match.groups()
                result_dict[f'{method}'] = f"{float(time):.3f}"
            for match in re.finditer(rate_pattern, output, re.M):
                method, rate = match.groups()
                result_dict[f'{method} GStencil/s'] = f"{float(rate):.3f}"
            outputs.append(result_dict)
    with open(args.out, 'w', newline='') as file:
        fieldnames = list(dict.fromkeys(k for o in outputs for k in o))
        writer = csv.DictWriter(file, fieldnames=fieldnames)
        writer.writeheader()
        writer.writerows(outputs)


if __name__ == "__main__":
    main()
