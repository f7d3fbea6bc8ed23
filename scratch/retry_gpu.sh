for i in 1 2 3 4 5 6 7 8; do
  /usr/local/graft/bin/gpurun --timeout 1200 -- 'bash scratch/prof_wgrad.sh' > gpurun_out/retry.log 2>&1
  if grep -q "status=ok" gpurun_out/retry.log; then exit 0; fi
  sleep 60
done
exit 1
