set -x
O=gpurun_out/r02u; mkdir -p $O
for V in 2 10; do for G in 2 8; do timeout 300 python tools/peer_local_check.py --views $V --peers $G >> $O/peer_local.jsonl 2>> $O/peer_local.err; done; done
cat $O/peer_local.jsonl; tail -5 $O/peer_local.err
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -4 $O/pytest_gpu.log
