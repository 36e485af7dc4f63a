#!/bin/bash
# Runs the executor-vs-reference comparison for several models, one process each.
mkdir -p gpurun_out
: > gpurun_out/exec_check.log
run() { echo "### $*" >> gpurun_out/exec_check.log; timeout 300 python tools/executor_check.py "$@" >> gpurun_out/exec_check.log 2>&1; echo "rc=$?" >> gpurun_out/exec_check.log; }
run --model trivial --batch 4
run --model resnet20 --data_name cifar10 --batch 8
run --model resnet20_v2 --data_name cifar10 --batch 8
run --model alexnet --data_name cifar10 --batch 8
run --model lenet --batch 4
run --model resnet50 --batch 4
run --model resnet50_v1.5 --batch 2
run --model resnet50_v2 --batch 2
run --model googlenet --batch 2
run --model inception3 --batch 2
grep -E "^###|RESULT|rc=|Error|error" gpurun_out/exec_check.log | cut -c1-600
