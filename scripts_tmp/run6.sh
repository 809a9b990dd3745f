cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r13_tests.log 2>&1; echo rc=$? >> gpurun_out/r13_tests.log
tail -5 gpurun_out/r13_tests.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --perceptual 0.1 --batch 16 --layer-report gpurun_out/r13_layers_vgg.txt > gpurun_out/r13_bench_vgg.log 2>&1
cut -c1-400 gpurun_out/r13_bench_vgg.log
grep -i "vgg\|conv_tc" gpurun_out/r13_layers_vgg.txt | head -40
