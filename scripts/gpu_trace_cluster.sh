#!/bin/bash
# phase timelines of the clustered solver: push mode (default), barrier.cluster with ghost rows, barrier.cluster without
F="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -DGBDPCG_TRACE -I gbd_pcg_b200/csrc -I include scripts/ubench/trace_cluster.cu"
for b in trace_cluster trace_cluster_nopush trace_cluster_noghost trace_cluster_c3; do echo "== $b"; timeout 20 scripts/ubench/$b 24; done
