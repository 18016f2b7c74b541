/* Mock of boost::crc_32_type (Boost.CRC, version unpinned by the reference).
 * Standard reflected CRC-32: poly 0xEDB88320 (0x04C11DB7 reflected), init 0xFFFFFFFF,
 * final xor 0xFFFFFFFF.  Used at /root/reference/lib/decode_impl.cc:372-374,451-453.
 * Test infrastructure only. */
#pragma once
#include <cstdint>
#include <cstddef>
namespace boost {
class crc_32_type {
    uint32_t d_rem;
    static const uint32_t* table() {
        static uint32_t t[256];
        static bool init = false;
        if (!init) {
            for (uint32_t i = 0; i < 256; i++) {
                uint32_t c = i;
                for (int k = 0; k < 8; k++) c = (c & 1) ? (0xEDB88320u ^ (c >> 1)) : (c >> 1);
                t[i] = c;
            }
            init = true;
        }
        return t;
    }
public:
    crc_32_type() : d_rem(0xFFFFFFFFu) {}
    void reset() { d_rem = 0xFFFFFFFFu; }
    void process_bytes(const void* buf, std::size_t n) {
        const uint8_t* p = static_cast<const uint8_t*>(buf);
        const uint32_t* t = table();
        for (std::size_t i = 0; i < n; i++) d_rem = t[(d_rem ^ p[i]) & 0xFF] ^ (d_rem >> 8);
    }
    uint32_t checksum() const { return d_rem ^ 0xFFFFFFFFu; }
};
}
