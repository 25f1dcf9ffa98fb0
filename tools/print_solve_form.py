import json,sys
d=json.load(open(sys.argv[1])); sf=d["solve_form"]
for k in ("reference_semantics","frontier_rules"): print(sys.argv[1][-28:], k, round(sf[k]["solves_per_s"]), round(sf[k]["wall_s"],3), round(sf[k]["us_per_lm_iteration_latency"],1), sf[k]["lm_iterations"])
