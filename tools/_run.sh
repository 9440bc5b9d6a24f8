for i in 1 2; do
for v in tools/libahk_head.bin tools/libahk_vH1.bin tools/libahk_vH2.bin tools/libahk_vH3.bin tools/libahk_vC.bin; do
echo -n "$v  "; AHK_SPREAD=0 AHK_LIB=$v python tools/launch_probe.py 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('stream %.2f graph %.2f' % (d['stream_gpu_us_per_step'], d['graph_gpu_us_per_step']))"
done
done
