import json,sys
d=json.load(open(sys.argv[1])); r=d["roofline"]
print(sys.argv[1], round(d["value"]), round(d["ms_per_step"],2), "contract", round(r["share_of_step"],3), "step", round(r["step_kernel_share"],3), round(r["achieved"]))
