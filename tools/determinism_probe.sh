#!/usr/bin/env bash
# Cross-process determinism probe (GPU box): N cold processes each run one full-size case through the public entry
# points and save inputs, integer outputs, maps and gradients; every run is compared BIT FOR BIT with the first.
#   bash tools/determinism_probe.sh C3v7 16      -> prints the keys that differ (nothing but "done N" when all agree)
c=${1:-C3v7}; n=${2:-8}
python tools/determinism_probe.py run $c /tmp/r_1.pt
for k in $(seq 2 $n); do
  python tools/determinism_probe.py run $c /tmp/r_k.pt
  python tools/determinism_probe.py cmp /tmp/r_1.pt /tmp/r_k.pt | grep -v "differ in: \[\]"
done
echo done $n
