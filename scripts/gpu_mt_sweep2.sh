# encoder time vs sub-tiles per tile (AT_TC_MT<N>) after the shortcut folds
for a in 1 2; do for b in 1 2; do for c in 1 2; do
  echo -n "MT32=$a MT64=$b MT128=$c: "
  AT_TC_MT32=$a AT_TC_MT64=$b AT_TC_MT128=$c timeout 100 python scripts/tc_time.py fp16 2048 2>&1 | tail -1
done; done; done
echo -n "defaults: "; timeout 100 python scripts/tc_time.py fp16 2048 2>&1 | tail -1
