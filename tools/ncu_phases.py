#!/usr/bin/env python
"""Per-phase warp-instruction table and opcode histogram of a MAP kernel from an ncu source-page dump
(`ncu -i x.ncu-rep --page source --csv --print-source cuda,sass`).  Every SASS instruction is counted once (by
address) and attributed to the source line ncu lists it under first; lines are mapped to phases by the function
they sit in (device functions) or by markers in the kernel body.

  python tools/ncu_phases.py dump.csv <units> [old|stream]      units = 512-byte rows the capture processed"""
import csv, re, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
dump, units = sys.argv[1], float(sys.argv[2])
kind = sys.argv[3] if len(sys.argv) > 3 else "old"
import subprocess
REV = os.environ.get("NCU_SRC_REV")          # the commit the captured library was built from (default: the working tree)
def _src(f):
    if REV:
        r = subprocess.run(["git", "-C", ROOT, "show", "%s:locust_b200/csrc/%s" % (REV, f)], capture_output=True, text=True)
        return r.stdout.split("\n") if r.returncode == 0 else []
    p = os.path.join(ROOT, "locust_b200", "csrc", f)
    return open(p).read().split("\n") if os.path.exists(p) else []
SRC = {f: _src(f) for f in ("map_kernel.cu", "map_stream_kernel.cu", "common.cuh", "long_keys.cuh", "smem_ptx.cuh")}
FUNC_PHASE = {  # device function -> phase
    "classify4": "A classify", "classify4_bal": "A classify", "classify16": "A classify", "classify16_bal": "A classify",
    "scan_step": "A starts/scan/list", "load_chunk": "A load", "mask_chunk": "A load", "ldg_stream": "A load",
    "process_batch": "B token: extract + probe", "smem_hit": "B token: extract + probe", "smem_bucket": "B token: extract + probe",
    "take_token": "B token: extract + probe", "extract_key16_ascii": "B token: extract + probe", "extract_key16_any": "B token: extract + probe",
    "key_selectors": "B token: extract + probe", "byte_perm": "B token: extract + probe", "shl_clamp": "B token: extract + probe",
    "drain_queue": "drain / insert (global table)", "global_insert": "drain / insert (global table)", "global_insert_at": "drain / insert (global table)", "smem_insert": "drain / insert (global table)",
    "insert_batches": "drain / insert (global table)", "insert_keys": "drain / insert (global table)", "global_probe": "drain / insert (global table)",
    "insert": "drain / insert (global table)", "hash_key128": "drain / insert (global table)", "mix64": "drain / insert (global table)",
    "slot_hash32": "drain / insert (global table)", "cas128": "drain / insert (global table)", "hand_over": "hand-over to inserter warps",
    "far_push": "long tokens", "far_token": "long tokens", "long_insert": "long tokens", "token_generic": "long tokens",
}
def func_of(f, ln):
    lines = SRC.get(f)
    if not lines: return None
    for i in range(min(ln, len(lines)) - 1, -1, -1):
        m = re.match(r"^(?:template.*)?(?:LC_HD|LC_D|__device__|__global__|static|inline)[^;(]*?(\w+)\s*\(", lines[i])
        if m: return m.group(1)
    return None
def body_phase(f, ln):
    lines = SRC[f]
    def first(marker, default=10**9):
        for i, l in enumerate(lines):
            if marker in l: return i + 1
        return default
    if f == "map_kernel.cu":
        a, b, fl = first("// A: stage bytes"), first("// B: one token per lane"), first("// Flush the CTA's combiner")
        loop = first("for (uint32_t i = 0; i < n_rows; ++i)")
        if ln >= fl: return "flush of the shared table"
        if ln >= b and ln < b + 6: return "B batch loop control"
        if ln >= a and ln < b: return "A starts/scan/list"
        if ln >= loop: return "A load"
        return "prologue"
    if f == "map_stream_kernel.cu":
        fl = first("// Flush the CTA's combiner")
        b = first("// --- B: two tokens per lane")
        a = first("// --- A: token starts")
        ins = first("inserter warp ====")
        if ln >= fl: return "flush of the shared table"
        if ln >= ins: return "drain / insert (global table)"
        if ln >= b: return "B batch loop control + queue"
        if ln >= a: return "A starts/scan/list"
        return "A load (bulk copy issue, wait, LDS)"
    return "other"
seen, cur, line, ops, phases = set(), None, None, {}, {}
tot = 0.0
for r in csv.reader(open(dump)):
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]; continue
    if len(r) > 10 and r[0].strip().isdigit() and r[2] == "-":
        line = int(r[0]); continue
    if len(r) > 10 and r[0] == "" and r[2].startswith("0x"):
        addr = r[2]
        if addr in seen: continue
        seen.add(addr)
        try: ie = int(r[7])
        except ValueError: continue
        f = cur
        fn = func_of(f, line) if f in SRC else None
        ph = FUNC_PHASE.get(fn)
        if ph is None:
            ph = body_phase(f, line) if f in ("map_kernel.cu", "map_stream_kernel.cu") else ("drain / insert (global table)" if f and "atomic" in f else "other (intrinsics headers)")
        phases[ph] = phases.get(ph, 0) + ie
        op = r[3].strip().split()[0]
        if op.startswith("@"): op = r[3].strip().split()[1]
        op = op.rstrip(";")
        ops[op] = ops.get(op, 0) + ie
        tot += ie
print("warp instructions per 512-byte row, by phase (total %.1f):" % (tot / units))
for k, v in sorted(phases.items(), key=lambda x: -x[1]):
    print("  %7.1f  %5.1f%%  %s" % (v / units, 100.0 * v / tot, k))
print("\nopcode histogram (executed warp instructions per 512-byte row, top 40):")
for k, v in sorted(ops.items(), key=lambda x: -x[1])[:40]:
    print("  %7.2f  %s" % (v / units, k))
for must in ("UBLKCP", "SYNCS", "LDG.E", "ATOM", "ATOMS", "REDG", "RED", "LDS", "STS"):
    hits = {k: v for k, v in ops.items() if k.startswith(must)}
    if hits: print("  [%s*] %s" % (must, ", ".join("%s %.2f" % (k, v / units) for k, v in sorted(hits.items(), key=lambda x: -x[1])[:6])))
