"""Embed the kernel template as a C++ string (split into chunks below the
compiler's string-literal limits)."""
import sys

src = open(sys.argv[1]).read()
chunks = [src[i:i + 8000] for i in range(0, len(src), 8000)]
with open(sys.argv[2], "w") as f:
    f.write("// generated from device_kernels.cuh by embed_source.py\n")
    f.write("#include <string>\nnamespace smeb200 {\n")
    f.write("static const std::string kSourceStorage = std::string()\n")
    for c in chunks:
        f.write('  + std::string(R"SMEBSRC(' + c + ')SMEBSRC")\n')
    f.write(";\nextern const char *const kDeviceKernelSource;\n")
    f.write("const char *const kDeviceKernelSource = kSourceStorage.c_str();\n}\n")
