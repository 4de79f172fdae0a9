# e2e leg of bench.py over chunk sizes and expansion threads (experiment script; results in gpurun_out/<tag>/)
O=gpurun_out/${1:-e2e_sweep}; mkdir -p $O
B="python bench.py --steps 40 --warmup 3 --no-secondary --no-cpu-baseline"
show() { python -c "
import json,sys; d=json.load(open(sys.argv[1])); e=d['e2e']; print(sys.argv[1], 'u8', round(e['ms_per_step'],3), 'blocking', round(e['ms_per_step_blocking'],3), 'f32', round(e['float32_mask_transport']['ms_per_step'],3), e['host_mask_bit_identical_to_float32_copy'])" $1; }
for c in 4 8 16; do for t in 2 4; do $B --e2e-chunk $c --e2e-mask-threads $t > $O/c${c}_t$t.json 2>/dev/null; show $O/c${c}_t$t.json; done; done
