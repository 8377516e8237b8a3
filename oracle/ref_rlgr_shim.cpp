// ref_rlgr_shim.cpp -- TEST INFRASTRUCTURE ONLY (see oracle/raht_oracle.c header).
//
// A C-ABI veneer over the reference's OWN entropy coder, compiled from the sources where they lie
// under /root/reference (python/PyRLGR/src/libs/rlgr/membuf.{cpp,h}) by oracle/Makefile into
// oracle/_ref/librlgr_ref.so.  No reference source is copied into this repo: this file only
// #includes "membuf.h" via -I at build time and calls the class' public methods the way the
// reference's pybind layer does (bindings.cpp:34-57; call sites encode_3dgs.py:229-245).
#include "membuf.h"
#include <cstring>
#include <vector>

extern "C" {

// rlgrWrite + close + get_buffer (encode_3dgs.py:229-233). Returns byte count; copies min(cap, n) bytes.
__attribute__((visibility("default")))
long long ref_rlgr_encode(const long long *seq, long long N, int flagSigned, unsigned char *out,
                          long long cap, long long *time_ns)
{
    membuf m;
    std::vector<int64_t> v(seq, seq + N);
    auto ns = m.rlgrWrite(v.data(), (size_t)N, (uint_least8_t)flagSigned);
    m.close();
    std::vector<uint8_t> b = m.get_buffer();
    if (time_ns) *time_ns = (long long)ns.count();
    long long n = (long long)b.size();
    if (out && n) std::memcpy(out, b.data(), (size_t)(n < cap ? n : cap));
    return n;
}

// membuf(buf).rlgrRead(N, flagSigned) (encode_3dgs.py:239-240).
__attribute__((visibility("default")))
void ref_rlgr_decode(const unsigned char *buf, long long nbytes, long long N, int flagSigned,
                     long long *seq, long long *time_ns)
{
    std::vector<uint8_t> b(buf, buf + nbytes);
    membuf m(b);
    std::vector<int64_t> v((size_t)N, 0);
    auto ns = m.rlgrRead(v.data(), (size_t)N, (uint_least8_t)flagSigned);
    if (time_ns) *time_ns = (long long)ns.count();
    for (long long i = 0; i < N; ++i) seq[i] = v[(size_t)i];
}

}  // extern "C"
