# A/B of the early programmatic-launch trigger on the bench step. Build the variant first (in triton_client_b200/csrc):
#   nvcc <flags of the Makefile> -DTCB_PDL_EARLY -c preprocess.cu yolo_post.cu, link with the other objects into
#   profiles/probes/libtcb200_early.so; then run this on the GPU box. Result of the round: 121.2 us (default) vs 122.6 us.
for v in "" profiles/probes/libtcb200_early.so "" profiles/probes/libtcb200_early.so; do
  TCB_LIB_PATH=$v python bench.py --no-cpu-baseline --steps 200 2>/dev/null > /tmp/o.json
  python -c "import json; d=json.load(open('/tmp/o.json')); print('lib=$v', round(d['value']), round(d['ms_per_step']*1000,2), d['kernel_ms'], round(d['roofline']['frac'],4), d['p50_latency']['pre_post_ms'])"
done
TCB_LIB_PATH=profiles/probes/libtcb200_early.so python -m pytest tests/test_gpu_yolo.py tests/test_gpu_pre2d.py tests/test_gpu_clients.py -x -q 2>&1 | tail -1
