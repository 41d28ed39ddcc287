#!/bin/bash
# round 2, call 8u: the C++ host test with the windowed Add selected through hnsw::Graph::SetAddWindow
mkdir -p gpurun_out
timeout 60 ./tests/cpp/host_test --gpu > gpurun_out/r8u_host_test.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/r8u_host_test.log
