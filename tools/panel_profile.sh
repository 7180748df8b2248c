#!/bin/bash
# builds tools/panel_profile (the panel kernel with its phase timers compiled in); run it on the GPU box: tools/panel_profile <n>
cd "$(dirname "$0")/.." && nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -DTP_PROFILE -Iiddiff_b200/csrc \
  tools/panel_profile.cu iddiff_b200/csrc/idb_eig_panel.cu -o tools/panel_profile
