#!/bin/bash
# lists local-memory (spill) accesses of the 7B persistent-kernel variant next to landmark instructions (tuning aid)
O=${1:-ollama-operator_b200/build/mega.o}
F=$(cuobjdump -elf $O 2>/dev/null | grep -o "_ZN2b2[A-Za-z0-9_]*mega_kernelILi2ELi4ELi15ELi1ELi128ELb0[A-Za-z0-9_]*S5_" | sort -u | head -1)
cuobjdump -sass -fun "$F" $O 2>/dev/null | grep -E "^\s+/\*[0-9a-f]{4,5}\*/" | sed -E 's/^\s+\/\*([0-9a-f]+)\*\/\s+/\1 /; s/\s*\/\*.*$//' > /tmp/mega_new.sass
grep -n "STL\|LDL\|MUFU.RCP\|BAR.SYNC\|IDP.4A.S8.S8 R[0-9]*, R[0-9]*, R[0-9]*, RZ\|CCTL.IVALL\|LDG.E.128.STRONG.GPU\|UBLKCP" /tmp/mega_new.sass | awk -F'[: ]' '{printf "%s:%s ", $1, $3} END{print ""}' | fold -w 200
