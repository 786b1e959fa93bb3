#!/bin/bash
mkdir -p gpurun_out
( time python -m pytest tests/test_annot_gpu.py -x -q ) > gpurun_out/r5e_annot_tests.log 2>&1
echo "rc=$?" >> gpurun_out/r5e_annot_tests.log
cd tests/golden && ( time ../../vt_b200/vtn annotate_indels annot_in.vcf -r annot_ref.fa -o /tmp/o.vcf ) > ../../gpurun_out/r5e_cli.log 2>&1; cd ../..
diff /tmp/o.vcf tests/golden/annot_out.vcf > gpurun_out/r5e_cli_diff.txt; echo "diff rc=$?" >> gpurun_out/r5e_cli.log
