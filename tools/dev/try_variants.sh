#!/bin/bash
# Rebuild the library with different tuning macros on the GPU box and time each (development aid).
N=${N:-5000}; M=${M:-1000}; K=${K:-4}
for v in "$@"; do
  MELISSA_NVCC_FLAGS="$v" python melissa_b200/build.py > /dev/null 2>&1 || { echo "build failed: $v"; continue; }
  echo "== $v"; python tools/dev/quick_time.py $N $M $K | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['kernels_ms'], 'pairs/s %.3e'%d['pairs_per_s'], 'lb', d['lb'])"
done
python melissa_b200/build.py > /dev/null 2>&1
