import json,sys
for f in sys.argv[1:]:
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f,"ERR",e); continue
    k=d["kernels"]
    print(f, "val %.1f e2e %.1f its %d pcgit %.0f"%(d["value"], d["e2e"]["value"], d["config"]["lm_iterations_per_solve"], d["config"]["pcg_iterations_per_solve"]), " ".join("%s=%.0f"%(n,k[n]["avg_us"]) for n in ("pcg","points","schur_blocks","img_chunks","back_substitute","coarse_setup") if n in k))
