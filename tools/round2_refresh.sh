O=gpurun_out/r2b; mkdir -p $O
for shape in "57344 8192" "8192 28672" "10240 8192" "8192 8192"; do set -- $shape; timeout 100 python tools/gemv_bench.py --N $1 --K $2 --b 64,128,256 2>&1 | grep q4_k; done > $O/gemm_tcgen05.txt
{ timeout 100 python tools/gemv_bench.py --type q6_k --N 32000 --K 8192 --b 1,4,16 --variant 0 2>&1 | grep q6_k
  timeout 100 python tools/gemv_bench.py --type q6_k --N 8192 --K 28672 --b 1,4,16 --variant 0 2>&1 | grep q6_k
  timeout 100 python tools/gemv_bench.py --type q5_k --N 8192 --K 28672 --b 1,4,16 --variant 0 2>&1 | grep q5_k
  timeout 100 python tools/gemv_bench.py --type q5_k --N 1024 --K 8192 --b 1,4 --variant 0 2>&1 | grep q5_k
  timeout 100 python tools/gemv_bench.py --type q3_k --N 57344 --K 8192 --b 1,4,16 --variant 0 2>&1 | grep q3_k
  timeout 100 python tools/gemv_bench.py --type q3_k --N 10240 --K 8192 --b 1,4 --variant 0 2>&1 | grep q3_k; } > $O/gemv_kquants.txt
timeout 200 python tools/ttft_bench.py > $O/ttft.txt 2>&1
cat $O/gemm_tcgen05.txt $O/gemv_kquants.txt $O/ttft.txt
