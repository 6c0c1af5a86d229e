"""Raw SASS evidence: split `cuobjdump -sass` of the shipped library per kernel.

usage: python scripts/dump_sass.py            (writes profiles/sass/*.sass + profiles/sass/README.md)
Every kernel gets a row of mnemonic counts in the README; the raw listing is written for the headline kernels (the full
dump is ~150 MB because the register-tiled statevector templates are instantiated for 1..10 qubits)."""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "qfedx_b200" / "lib" / "libqfedx_b200.so"
OUT = ROOT / "profiles" / "sass"
INTEREST = ["LDGMC", "UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "LDGSTS", "HMMA",
            "FFMA", "SHFL", "ATOMS", "ATOMG", "ATOM", "REDG", "RED", "MEMBAR", "CCTL", "BAR", "LDS", "STS", "LDG", "STG"]
# kernels whose raw listing is kept (regex on the demangled name)
KEEP = [r"tinycnn_round_kernel", r"tinycnn_eval_conv_kernel", r"conv2_(fwd|dgrad|wgrad)_umma_train", r"unitary_gemm_kernel<128>",
        r"sv_forward_kernel<8>", r"sv_adjoint_kernel<8>", r"vqc_local_train_kernel<8, 5>",
        r"sv_large_pass_v2_kernel", r"sv_large_adjoint_pass_kernel",
        r"sv_large_expz_masks_kernel", r"sv_global_swap_fused_kernel<3>", r"fedavg_fused_kernel<4, 2, (true|false), (true|false)>",
        r"p2p_probe_kernel<4>", r"readout_noise_kernel"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    txt = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True).stdout
    OUT.mkdir(parents=True, exist_ok=True)
    for old in OUT.glob("*.sass"):
        old.unlink()
    parts = re.split(r"\n\s*Function : ", txt)
    funcs = {}
    for part in parts[1:]:
        name, _, body = part.partition("\n")
        funcs[name.strip()] = body
    dm = demangle(list(funcs))
    rows = []
    for mangled, body in sorted(funcs.items(), key=lambda kv: dm[kv[0]]):
        name = dm[mangled]
        short = re.sub(r"\(.*", "", name).replace("void ", "")
        instr = re.findall(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", body, flags=re.M)
        counts = collections.Counter()
        for ins in instr:
            base = ins.split(".")[0]
            key = ins if (base in ("SYNCS", "MEMBAR") or ins.endswith(".SYS") or ".SYS." in ins) else base
            if any(key.startswith(i) for i in INTEREST):
                if ".SYS" in ins and base in ("LDG", "STG", "LD", "ST", "ATOMG", "REDG"):
                    key = base + ".SYS"
                counts[key] += 1
        kept = ""
        if any(re.search(k, name) for k in KEEP):
            fn = re.sub(r"[^A-Za-z0-9_]+", "_", short).strip("_")[:120] + ".sass"
            # instruction text only: the hex encoding columns triple the size and add nothing a reader checks
            slim = re.sub(r"[ \t]*/\* 0x[0-9a-f]{16} \*/", "", body)
            slim = "\n".join(l.rstrip() for l in slim.splitlines() if l.strip())
            (OUT / fn).write_text(f"// {name}\n// cuobjdump -sass {LIB.name} (sm_100a), encoding columns stripped\n" + slim + "\n")
            kept = fn
        notable = ", ".join(f"{k} {v}" for k, v in sorted(counts.items()) if k.split(".")[0] not in ("LDS", "STS", "LDG", "STG", "BAR") or ".SYS" in k)
        rows.append((short, len(instr), notable, kept))
    lines = ["# SASS evidence per kernel (`cuobjdump -sass qfedx_b200/lib/libqfedx_b200.so`, sm_100a)", "",
             "Regenerate with `python scripts/dump_sass.py` after a build.  `UTC*MMA` = tcgen05.mma, `LDTM` = tcgen05.ld, `UTMALDG` = TMA",
             "load, `SYNCS.*` = mbarrier operations, `LDGSTS` = cp.async, `*.SYS` = system-scope loads / stores / fences used on peer",
             "memory, `UTCBAR` = tcgen05.commit.  Raw listings are kept for the headline kernels (last column); the full dump is",
             "~150 MB because the register-tiled statevector templates are instantiated for 1..10 qubits.", "",
             "| kernel | instructions | notable mnemonics | raw listing |", "|---|---|---|---|"]
    for short, n, notable, kept in rows:
        lines.append(f"| `{short}` | {n} | {notable} | {kept} |")
    (OUT / "README.md").write_text("\n".join(lines) + "\n")
    print(f"{len(rows)} kernels, {sum(1 for r in rows if r[3])} raw listings in {OUT}")


if __name__ == "__main__":
    sys.exit(main())
