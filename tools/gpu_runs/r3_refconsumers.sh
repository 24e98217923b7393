#!/bin/bash
# round 3: the reference's OWN consumers (python/dftd4 through its CFFI module, test/api/example.c) evaluating on a GPU.
# The GPU box has no /root/reference, so the seven files the test needs are staged -- untracked, git-ignored -- into
# gpurun_stage/reference/ of the tree gpurun pushes, and removed again afterwards:
#   here:        bash tools/gpu_runs/r3_refconsumers.sh stage
#                gpurun -- 'bash tools/gpu_runs/r3_refconsumers.sh run'
#                bash tools/gpu_runs/r3_refconsumers.sh clean
set -e
STAGE=gpurun_stage/reference
case "$1" in
  stage)
    REF=${DFTD4_REFERENCE:-/root/reference}
    mkdir -p $STAGE/python/dftd4 $STAGE/test/api
    for f in __init__ library interface test_library test_interface; do cp $REF/python/dftd4/$f.py $STAGE/python/dftd4/; done
    cp $REF/python/ffibuilder.py $STAGE/python/
    cp $REF/test/api/example.c $STAGE/test/api/
    find $STAGE -type f | sort ;;
  run)
    mkdir -p gpurun_out
    DFTD4_REFERENCE=$PWD/$STAGE timeout 1500 python -m pytest tests/test_reference_consumers.py -v -rs \
      > gpurun_out/r3_reference_consumers.log 2>&1 || true
    tail -12 gpurun_out/r3_reference_consumers.log ;;
  clean)
    rm -rf gpurun_stage ;;
  *) echo "usage: $0 stage|run|clean"; exit 2 ;;
esac
