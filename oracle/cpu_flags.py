"""Print the x86 ISA flags that `gcc -march=native` enabled on the build host, as /proc/cpuinfo names.

Written next to oracle/_ref/native/libggml-cpu.so so that the GPU box can decide whether that variant is
safe to load there (see oracle/refenv.py).  Test infrastructure only.
"""
import re
import subprocess

# gcc macro -> /proc/cpuinfo flag
MACROS = {
    "__AVX__": "avx", "__AVX2__": "avx2", "__FMA__": "fma", "__F16C__": "f16c", "__BMI2__": "bmi2",
    "__AVX512F__": "avx512f", "__AVX512BW__": "avx512bw", "__AVX512DQ__": "avx512dq", "__AVX512VL__": "avx512vl",
    "__AVX512CD__": "avx512cd", "__AVX512VBMI__": "avx512vbmi", "__AVX512VBMI2__": "avx512_vbmi2",
    "__AVX512VNNI__": "avx512_vnni", "__AVX512BF16__": "avx512_bf16", "__AVX512FP16__": "avx512_fp16",
    "__AVX512IFMA__": "avx512ifma", "__AVX512BITALG__": "avx512_bitalg", "__AVX512VPOPCNTDQ__": "avx512_vpopcntdq",
    "__AVXVNNI__": "avx_vnni", "__AMX_TILE__": "amx_tile", "__AMX_INT8__": "amx_int8", "__AMX_BF16__": "amx_bf16",
    "__GFNI__": "gfni", "__VAES__": "vaes", "__VPCLMULQDQ__": "vpclmulqdq", "__SHA__": "sha_ni",
    "__MOVBE__": "movbe", "__ADX__": "adx", "__LZCNT__": "abm", "__POPCNT__": "popcnt", "__BMI__": "bmi1",
}


def main() -> None:
    out = subprocess.run(["gcc", "-march=native", "-dM", "-E", "-"], input="", capture_output=True, text=True).stdout
    defined = set(re.findall(r"#define (\w+) ", out))
    print(" ".join(sorted(v for k, v in MACROS.items() if k in defined)))


if __name__ == "__main__":
    main()
