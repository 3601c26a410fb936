#!/bin/bash
# per-kernel SASS size (instructions x 16 bytes), registers, stack frame of the library's kernels
for o in build/obj/avr_lib.cu.o build/obj/avr_probe.cu.o; do
  cuobjdump -sass $o | awk '/Function :/{if(name!="")printf "%8d instr %8.1f KB  %s\n", n, n*16/1024, name; name=$3; n=0} /^ +\/\*[0-9a-f]+\*\/ +[A-Z@]/{n++} END{printf "%8d instr %8.1f KB  %s\n", n, n*16/1024, name}' | c++filt | sed 's/(anonymous namespace):://g; s/((anonymous namespace)::[A-Za-z]*)//'
  cuobjdump -res-usage $o 2>/dev/null | grep -A1 "Function" | paste - - | sed 's/Fatbin.*//' | awk '{print $2, $4, $5, $6}' | c++filt | sed 's/(anonymous namespace):://g; s/((anonymous namespace)::[A-Za-z]*)//' | cut -c1-160
done
