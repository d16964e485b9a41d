#!/bin/bash
# resident-block sweep of the packet kernel (development aid)
for mb in 4 6 7 8; do
  make -C cudaraytracer_b200/csrc clean > /dev/null
  make -C cudaraytracer_b200/csrc -j8 EXTRA="-DCRT_PACKET_MIN_BLOCKS=$mb" > /dev/null 2>&1
  echo "== packet min_blocks $mb: $(cuobjdump -res-usage cudaraytracer_b200/csrc/build/render.o 2>/dev/null | grep -A1 'k_trace_packetILb0ELb0' | grep -o 'REG:[0-9]*')"
  python scripts/b0_probe.py 2>&1 | grep "mode 1"
done
make -C cudaraytracer_b200/csrc clean > /dev/null
make -C cudaraytracer_b200/csrc -j8 > /dev/null 2>&1
