export GRIDB200_GRAM_NOSYM=1
b() { timeout 120 python tools/gram_bench.py 6250 1000000 200 3; }
echo "== base"; b
for w in 64 128; do echo "== window $w"; GRIDB200_GRAM_WINDOW_MB=$w b; done
echo "== nopace"; GRIDB200_GRAM_NOPACE=1 b
for g in 2 6 8 12; do echo "== group_m $g"; GRIDB200_GRAM_GROUP_M=$g b; done
echo "== group_m 8 window 64";  GRIDB200_GRAM_GROUP_M=8 GRIDB200_GRAM_WINDOW_MB=64 b
echo "== static group 8"; GRIDB200_GRAM_SCHED=static GRIDB200_GRAM_GROUP_M=8 b
