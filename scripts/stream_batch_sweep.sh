# How many files of a packed wave cross the link interleaved (scheduler.drain_wave): big pieces, and the slab workload.
# TSB200_STREAM_BATCH=n: n slab-sized files per batch (large files count half); unset: half of the writers.
for B in ${BATCHES:-default 8 16}; do
  if [ "$B" = default ]; then unset TSB200_STREAM_BATCH; else export TSB200_STREAM_BATCH=$B; fi
  echo "== TSB200_STREAM_BATCH=$B: 20 x 400 MiB"; TSB200_TRACE_WRITES=0 PIECES=20 PIECE_MIB=400 python scripts/big_piece_take_trace.py 2>&1 | grep trial
  echo "== TSB200_STREAM_BATCH=$B: 16 x 512 MiB"; TSB200_TRACE_WRITES=0 PIECES=16 PIECE_MIB=512 python scripts/big_piece_take_trace.py 2>&1 | grep trial
  echo "== TSB200_STREAM_BATCH=$B: 40 x 200 MiB"; TSB200_TRACE_WRITES=0 PIECES=40 PIECE_MIB=200 python scripts/big_piece_take_trace.py 2>&1 | grep trial
  echo "== TSB200_STREAM_BATCH=$B: C2 state, e2e take"; SWEEP=0 python scripts/e2e_take_sweep.py 2>&1 | grep take_ms | cut -c1-200
done
