import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]);k=d["ksp"]
print("asm ms/step",d["ms_per_step"],"K frac",d["roofline"]["frac"])
print("csr it",k["csr_stream"]["ms_per_iteration"],"spmv",k["csr_stream"]["spmv"]["ms"])
print("eop it",k["element_operator"]["ms_per_iteration"],"mult",k["element_operator"]["spmv"]["ms"])
