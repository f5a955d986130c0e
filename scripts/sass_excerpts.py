"""SASS evidence for the claims DESIGN.md makes about the hot kernels: which instructions move the
data (UBLKCP = TMA-engine bulk copy global->shared, SYNCS = mbarrier, LDGSTS = cp.async, STG.E.NA.128 =
16-byte no-allocate stores), which do the math (MUFU.EX2, FFMA), and that no tensor-core or tensor-map
instruction is present (none is wanted: nothing on this path is a dense contraction).
Runs anywhere nvcc's cuobjdump is installed:  python scripts/sass_excerpts.py > profiles/r02_sass_excerpts.txt"""
import collections, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "areal_vllm_b200", "libb200rl.so")
WANT = [("logprob_fwd_kernel<0, 24, 2, 1, 1>", "logprob_fwd_kernelILi0ELi24ELi2ELb1ELb1EE"),
        ("logprob_bwd_kernel<0, 0>", "logprob_bwd_kernelILi0ELb0EE"),
        ("grpo_fused_kernel<0, 24, 2, 1, 0, pair=1> (clusters of two CTAs; STAS = st.async into distributed shared memory)",
         "grpo_fused_kernelILi0ELi24ELi2ELb1ELb0ELb1EE"),
        ("grpo_fused_kernel<0, 24, 2, 1, 0, pair=0>", "grpo_fused_kernelILi0ELi24ELi2ELb1ELb0ELb0EE"),
        ("gae_pieces_kernel", "gae_pieces_kernel"),
        ("gae_serial_kernel", "gae_serial_kernelE"),
        ("gae_serial_slots_kernel", "gae_serial_slots_kernel"),
        ("gae_serial_relay_kernel", "gae_serial_relay_kernel"),
        ("gae_serial_team_kernel (chain lane: per four tokens 4 FFMA + LDS.128 + STS.128)", "gae_serial_team_kernel"),
        ("interval_copy_kernel (first instance)", "interval_copy_kernel"),
        ("interval_move_kernel (first instance)", "interval_move_kernel")]
KEY = re.compile(r"\b(UBLKCP|SYNCS|LDGSTS|LDGDEPBAR|MUFU\.\w+|STG\.E\.\w*128|STG\.E\.NA\.128|LDG\.E\.\w*128|LDS\.128|STS\.128|"
                 r"STAS\.?\w*|UCGABAR_\w+|MEMBAR\.\w+\.\w+|UTMALDG|UTMASTG|UTCMMA|UTCHMMA|HMMA|LDTM|STTM|F2FP\.\w+|VHMNMX|SHFL\.\w+|ATOM\w*|RED\w*)")

sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
funcs, cur = collections.OrderedDict(), None
for ln in sass.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = m.group(1)
        funcs[cur] = []
    elif cur and re.match(r"\s+/\*[0-9a-f]{4}\*/", ln):
        funcs[cur].append(re.sub(r"/\* 0x[0-9a-f]+ \*/", "", ln).strip())
print(f"# cuobjdump -sass areal_vllm_b200/libb200rl.so ({len(funcs)} kernels, sm_100a); per kernel: instruction count, histogram of the")
print("# data-movement / special-function mnemonics, and the first occurrence of each as an excerpt.")
tot = collections.Counter()
for body in funcs.values():
    for ins in body:
        for k in KEY.findall(ins):
            tot[k] += 1
print("# whole library:", dict(tot.most_common()))
print("# tensor-core / tensor-map instructions in the whole library (expected none):",
      {k: v for k, v in tot.items() if k.startswith(("UTMA", "UTC", "HMMA", "LDTM", "STTM"))} or "none")
for title, pat in WANT:
    name = next((n for n in funcs if pat in n), None)
    if not name:
        print(f"\n== {title}: not found")
        continue
    body = funcs[name]
    hist, first = collections.Counter(), collections.OrderedDict()
    for ins in body:
        for k in KEY.findall(ins):
            hist[k] += 1
            first.setdefault(k, ins)
    print(f"\n== {title}  [{name[:100]}]  {len(body)} instructions")
    print("   " + ", ".join(f"{k} x{v}" for k, v in hist.most_common()))
    for k, ins in first.items():
        print("     " + ins[:150])
