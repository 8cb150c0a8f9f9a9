# round-2 evidence (one gpurun call each; ncu and compute-sanitizer never in the same call -- B200_PROFILING.md):
#   bash tools/round_end_evidence.sh tests      GPU test-suite
#   bash tools/round_end_evidence.sh bench      both bench arms + forward / training kernel splits
#   bash tools/round_end_evidence.sh ncu        launch list + ncu --set full of the forward and backward kernels
#   bash tools/round_end_evidence.sh sanitize   compute-sanitizer memcheck of the search / emit / segment-sum kernels
set -u
O=gpurun_out
case "${1:-tests}" in
tests)
  python -m pytest tests -m gpu -q --timeout 900 2>&1 | tail -15 > $O/r02_tests.log; cat $O/r02_tests.log ;;
bench)
  ( time timeout 1500 python bench.py > $O/r02_bench.json 2> $O/r02_bench.err ) 2>&1 | grep real; tail -3 $O/r02_bench.err
  ( time timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/r02_ref.json 2> $O/r02_ref.err ) 2>&1 | grep real
  for c in 1 2 4 8; do LIPVQ_FWD_CHUNKS=$c python tools/fwd_probe.py; done > $O/r02_fwd_chunks.jsonl 2>&1; cat $O/r02_fwd_chunks.jsonl
  python tools/train_probe.py | tail -25 > $O/r02_train_kernels.log; cat $O/r02_train_kernels.log ;;
ncu)
  python bench.py --steps 4 --warmup 3 --no-cpu --skip parity,latency,config4,config5,vqvae > $O/r02_plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file $O/r02_launches.csv python bench.py --steps 4 --warmup 3 --no-cpu --skip parity,latency,config4,config5,vqvae > $O/r02_ncu.log 2>&1
  LIPVQ_FWD_CHUNKS=1 python tools/fwd_probe.py > $O/r02_plain2.log 2>&1 &&
  LIPVQ_FWD_CHUNKS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"enc16_fused_kernel|vq_tc_kernel|quantize_emit_kernel" --launch-skip 9 -c 3 -o $O/r02_full -f python tools/fwd_probe.py > $O/r02_ncufull.log 2>&1; tail -2 $O/r02_ncufull.log
  python tools/train_probe.py > $O/r02_plain3.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"umma_kernel|seg0_kernel" --launch-skip 10 -c 6 -o $O/r02_full_bwd -f python tools/train_probe.py > $O/r02_ncufull_bwd.log 2>&1; tail -2 $O/r02_ncufull_bwd.log ;;
sanitize)
  python -m pytest tests/test_gpu_scale_parity.py -m gpu -q -k "small_forward_kernel_shapes or segment_sum_deterministic and 5000 or tiny_dims" > $O/r02_plain4.log 2>&1 &&
  timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_scale_parity.py tests/test_gpu_parity.py -m gpu -q -x -k "small_forward_kernel_shapes or (segment_sum and 5000) or tiny_dims or search_vs_oracle_shapes or fused_paths_ragged" > $O/r02_memcheck.log 2>&1; echo rc=$?; tail -15 $O/r02_memcheck.log ;;
esac
