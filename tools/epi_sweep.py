import sys, os
sys.path.insert(0, os.getcwd())
import __graft_entry__ as ge
pkg = ge.load_package()
M=131072
for wt in (1,2):
    for epi in (0,1,3,9,11):
        ms, tf = pkg.bench_gemm(M, 3072, 768, wt, epi, reps=10)
        print("wt",wt,"epi",epi,round(ms,4),round(tf,1), flush=True)
    for epi in (1,5):
        ms, tf = pkg.bench_gemm(M, 768, 768, wt, epi, reps=10)
        print("o wt",wt,"epi",epi,round(ms,4),round(tf,1), flush=True)
