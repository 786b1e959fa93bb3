#!/bin/bash
# tool library split + VariantManip facade + vtn_variant_props
mkdir -p gpurun_out
python -m pytest tests/test_variant_manip_gpu.py tests/test_cli_gpu.py tests/test_annot_gpu.py tests/test_ref_lazy_gpu.py -x -q -m gpu > gpurun_out/r5m_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r5m_tests.txt
