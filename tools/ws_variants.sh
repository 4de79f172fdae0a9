# A/B of the persistent kernel's shared-memory split (band capacity WS_RMAX x stage ring WS_STAGES): rebuilds the library on the box
# per variant and times the 16-frame and one-frame cfg2 call (tools/ws_check.py --time-only).   gpurun -- 'bash tools/ws_variants.sh tag "512 4" "384 6" ...'
TAG=$1; shift; O=gpurun_out/$TAG; mkdir -p $O
for v in "$@"; do set -- $v; R=$1; S=$2
  if RMVD_BUILD_DEFS="WS_RMAX=$R WS_STAGES=$S" python -c "from robustmvd_b200 import _build; _build.build(force=True)" > $O/build_${R}_$S.log 2>&1; then
    python tools/ws_check.py --time-only 2>/dev/null | grep timing | sed "s/^/RMAX=$R STAGES=$S /" | cut -c1-200 | tee -a $O/variants.log
  else echo "RMAX=$R STAGES=$S build failed" | tee -a $O/variants.log; tail -3 $O/build_${R}_$S.log; fi
done
