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
    std::vector<unsigned long long> tok((size_t)out_len / 3 + 2);
    return svtk::bz_inflate(w.data(), (unsigned long long)in_beg, (unsigned long long)in_end, out, out_len, T, tok.data());
}

// CRC-32 of data[0..n) computed the way the kernel does: `slices` slice CRCs merged with the x^(8n) identity
extern "C" unsigned hostsim_crc32(const uint8_t *data, unsigned n, unsigned slices) {
    uint32_t tab[256];
    for (uint32_t i = 0; i < 256; ++i) tab[i] = svtk::bz_crc_table_entry(i);
    const unsigned per = (n + slices - 1) / slices;
    uint32_t crc = 0;
    for (unsigned s = 0; s < slices; ++s) {
        const unsigned long long b = (unsigned long long)s * per;
        const unsigned len = b >= n ? 0u : (n - (unsigned)b < per ? n - (unsigned)b : per);
        if (len == 0 && s > 0) continue;                 // empty slices contribute nothing (an all-empty message: slice 0 gives 0)
        crc ^= svtk::bz_crc_share(svtk::bz_crc_bytes(tab, data + b, len), s, per, n);
    }
    return crc;
}
