#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "batch_process or reference_api or load_audio or sharded or wav" > gpurun_out/pytest_bp.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_bp.log
python tools/bench_batch_process.py 2000 10 2>/dev/null | tail -1 | tee gpurun_out/batch_process_files.jsonl
AFX_INGEST_CHUNK_SAMPLES=$((32<<20)) python tools/bench_batch_process.py 2000 10 2>/dev/null | tail -1
AFX_INGEST_CHUNK_SAMPLES=$((64<<20)) python tools/bench_batch_process.py 2000 10 2>/dev/null | tail -1
python tools/bench_batch_process.py 6000 10 2>/dev/null | tail -1
