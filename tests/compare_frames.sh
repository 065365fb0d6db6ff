# benchmarks-cores/compare.py (the reference's own regression protocol, threshold 1e-8) on the
# frames tests/bridge_perf.py brought back from the GPU box: <dir>/frames/{ref,gpu}/<model>.save.N
# Runs where /root/reference exists (the CPU container).
#   bash tests/compare_frames.sh gpurun_out/r2/bridge
d=${1:-gpurun_out/r2/bridge}
for f in $d/frames/ref/*.save.*; do
  b=$(basename $f); model=${b%%.save.*}; frame=$((10#${b##*.}))
  [ $frame -eq 0 ] && continue      # a stand-in from an earlier run of this script (see below)
  # Dynearthsol.py probes the FIRST frame listed in <model>.info for the file format; only the
  # last frame was brought back, so it stands in for frame 0 (never compared)
  for side in ref gpu; do [ -e $d/frames/$side/$model.save.000000 ] || cp $d/frames/$side/$b $d/frames/$side/$model.save.000000; done
  echo "== $model frame $frame"
  (cd /root/reference/benchmarks-cores && python compare.py $OLDPWD/$d/frames/ref/$model $OLDPWD/$d/frames/gpu/$model $frame)
  echo "compare.py exit code: $?"
done
