#!/bin/bash
# Executor-vs-reference comparison for the wider zoo (padded channels, depthwise, ...).
mkdir -p gpurun_out
: > gpurun_out/exec_check2.log
run() { echo "### $*" >> gpurun_out/exec_check2.log; timeout 300 python tools/executor_check.py "$@" >> gpurun_out/exec_check2.log 2>&1; echo "rc=$?" >> gpurun_out/exec_check2.log; }
for m in "$@"; do
  case $m in
    densenet*|resnet20*|nasnet_cifar) run --model ${m%_cifar} --data_name cifar10 --batch 8 ;;
    ssd300) run --model ssd300 --data_name coco --batch 2 ;;
    *) run --model $m --batch 2 ;;
  esac
done
grep -E "^###|RESULT|rc=|Error|error" gpurun_out/exec_check2.log | cut -c1-700
