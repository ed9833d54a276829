# e2e (host buffers) against the number of upload chunks; then the host-buffer parity test
cd $GRAFT_REPO_ROOT
for c in 4 2 3; do
  BAMCL_HOST_CHUNKS=$c python bench.py --steps 20 --warmup 3 --legs none --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readline())
print('chunks $c: e2e %.3f ms (update half %.3f) resident %.3f hash %s' % (d['e2e']['ms_per_step'], d['e2e']['update_ms'], d['ms_per_step'], d['config']['result_hash']))"
done > gpurun_out/e2e_chunks2.txt 2>&1
python -m pytest tests/test_gpu_resample.py -m gpu -q -k "host_buffer" 2>&1 | tail -2 >> gpurun_out/e2e_chunks2.txt
