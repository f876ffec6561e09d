import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
from sparrowhawk_b200 import AssemblyHelper, synth
from parity import assert_parity, run_oracle
k = int(sys.argv[1])
seq, qual = synth.reads(61, 60_000, 25.0, 0.01)
h = AssemblyHelper.new(k, False, 2, 20, 0, False, False, False, False)
o = run_oracle(seq, qual, k, 2, 20)
h.load_reads(seq, qual); h.run(); assert_parity(h, o, check_table=False)
h.run(); assert h.stats()["device_planned"] == 1; assert_parity(h, o, check_table=False)
os.environ["X"] = "1"
h.preprocess_reads(seq, qual); h.assemble(); assert_parity(h, o, check_table=False)
print("case ok", k)
