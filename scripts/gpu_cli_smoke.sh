#!/bin/bash
# user-facing entry points on one GPU: PBT CLI (file exploit default, p2p, new engine options), stand-alone run loop
O=gpurun_out/r9; mkdir -p $O; export PYTHONPATH=$PWD; cd $O
run() { echo "== $*"; timeout 300 "$@" > cli.log 2>&1; echo "rc=$?"; tail -4 cli.log | cut -c1-220; }
run python -m distributedtf_b200 4 --model cifar10 --rounds 2 --max-train-steps 5 --eval-batches 2 --resnet-size 8 --save-root s1 --results-file r1.txt
run python -m distributedtf_b200 4 --model cifar10 --rounds 2 --max-train-steps 5 --eval-batches 2 --resnet-size 11 --resnet-bottleneck --dtype tf32 --exploit-path p2p --engine native --save-root s2 --results-file r2.txt
run python -m distributedtf_b200 4 --model mnist --rounds 2 --max-train-steps 5 --eval-batches 2 --exploit-path p2p --workers 2 --save-root s3 --results-file r3.txt
run python -m distributedtf_b200.models.resnet_run --model_dir m1 --use_synthetic_data --max_train_steps 5 --train_epochs 2 --resnet_size 8 --export_dir e1
run python -m distributedtf_b200.models.resnet_run --model_dir m2 --use_synthetic_data --max_train_steps 5 --train_epochs 1 --resnet_size 11 --resnet_bottleneck --resnet_version 1 --dtype fp32
ls s1 s2/model_0 e1 2>&1 | head -20
