#pragma once
#include <cstdint>
#include <pmt/pmt.h>
namespace gr {
struct tag_t {
    uint64_t offset;
    pmt::pmt_t key;
    pmt::pmt_t value;
    pmt::pmt_t srcid;
    tag_t() : offset(0), key(pmt::PMT_NIL), value(pmt::PMT_NIL), srcid(pmt::PMT_NIL) {}
};
} // namespace gr
