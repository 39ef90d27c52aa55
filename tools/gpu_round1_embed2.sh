#!/bin/bash
# Embedder wired into the CLIs: embed tests (incl. FASTA/FASTQ-only CLI run) + the existing CLI test.
set -x
mkdir -p gpurun_out
timeout 80 python -m pytest tests/test_gpu_zz_embed.py "tests/test_gpu_parity.py::test_cli_hnswpq_index_then_pipeline_config1" -q > gpurun_out/pytest_embed2.log 2>&1
echo "pytest rc=$?"
tail -40 gpurun_out/pytest_embed2.log
