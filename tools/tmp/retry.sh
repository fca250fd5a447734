# usage: retry.sh <timeout> <script>; retries while the pod answers busy (rc 3)
for i in 1 2 3 4 5 6 7 8; do
  /usr/local/graft/bin/gpurun ${GPUS:+--gpus $GPUS} --timeout $1 -- "bash $2" > /tmp/gpurun_last.txt 2>&1
  if ! grep -q "status=transient" /tmp/gpurun_last.txt; then break; fi
  sleep 120
done
tail -${3:-20} /tmp/gpurun_last.txt
