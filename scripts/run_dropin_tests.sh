#!/bin/bash
# Runs the reference's own Catch test binaries, linked against the GPU-backed
# pt::pll::Partition (oracle/Makefile.gpu), on the GPU box.
G=$PWD/oracle/_ref/gpu
mkdir -p gpurun_out
( cd $G/testdata/ptpll && timeout 600 $G/test_ptpll_gpu "[partition],[optimization],[util],[model_getter]" -d yes ) > gpurun_out/dropin_ptpll.log 2>&1
echo "test_ptpll_gpu rc=$?"; tail -5 gpurun_out/dropin_ptpll.log
( cd $G/testdata/ptw && timeout 1500 $G/test_ptw_gpu -d yes ) > gpurun_out/dropin_ptw.log 2>&1
echo "test_ptw_gpu rc=$?"; tail -40 gpurun_out/dropin_ptw.log
