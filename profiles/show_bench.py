import json,sys
d=json.load(open(sys.argv[1])); print(sys.argv[1], round(d["value"],1),"Mcells/s", round(d["ms_per_step"],3),"ms/step")
print("  "+"  ".join(f'{k}={v["ms_per_step"]:.3f}' for k,v in d["per_kernel"].items() if v["ms_per_step"]>0.04))
