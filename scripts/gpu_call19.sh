# GPU call 19: the tail of the W4 launch: ticks per circuit, then ncu of the two slowest circuits alone
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python scripts/w4_slowest.py rank 2500 > gpurun_out/c19_rank.txt 2>&1; cat gpurun_out/c19_rank.txt | cut -c1-250
two=$(grep "^circuit " gpurun_out/c19_rank.txt | head -2 | awk '{print $2}' | tr '\n' ' ')
echo "slowest: $two"
python scripts/w4_slowest.py only $two
ncu --set full --import-source on --clock-control none -k regex:dusp_render_tile -s 1 -c 1 -f -o gpurun_out/c19_slowest python scripts/w4_slowest.py only $two > gpurun_out/c19_ncu.log 2>&1
ls -la gpurun_out/c19*
