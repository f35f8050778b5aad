#!/bin/bash
# Example single run — the reference's run.sh (bsub ... ./build/bin/stencil_main ...) with
# the LSF/Sunway launcher removed and the flags brought up to the current CLI
# (reference run.sh:3 still used the stale `-m <size>`; -s is the size, -m the methods).
# On the build container (no GPU) prefix with: /usr/local/graft/bin/gpurun --
python -c 'import __graft_entry__ as g; g.build()' &&
./stencil_b200/host/build/bin/stencil_main -s 400 -b 50 -i 1000 -r 1 -m CUDA "$@"
