// The DEFLATE decoder of svtrek_b200/csrc/svtk_bgzf_core.cuh compiled as plain C++ (one "lane"), for CPU tests against zlib.
#include <cstring>
#include <vector>
#include "../../svtrek_b200/csrc/svtk_bgzf_core.cuh"

extern "C" int hostsim_inflate(const uint8_t *comp, long long n, long long in_beg, long long in_end, uint8_t *out, unsigned out_len) {
    std::vector<uint32_t> w((size_t)(n + svtk::BZ_OVERREAD) / 4 + 4, 0u);
    memcpy(w.data(), comp, (size_t)n);
    static uint16_t lit[1 << svtk::BZ_LBITS], dist[1 << svtk::BZ_DBITS];
    svtk::BzTables T;
    T.lit = lit; T.dist = dist;
    return svtk::bz_inflate(w.data(), (unsigned long long)in_beg, (unsigned long long)in_end, out, out_len, T);
}
