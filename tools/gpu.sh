#!/bin/bash
# build the in-tree CUDA library if stale, then run a command on the GPU box:  tools/gpu.sh [--timeout S] -- 'cmd'
cd "$(dirname "$0")/.."
python -c "import __graft_entry__ as g; g.build()" >/dev/null 2>&1 || { echo "build failed"; python -c "import __graft_entry__ as g; g.build()" 2>&1 | grep -E "error" | head; exit 1; }
exec /usr/local/graft/bin/gpurun "$@"
