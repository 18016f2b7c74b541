/* Mock of <gnuradio/io_signature.h>.  Test infrastructure only. */
#pragma once
#include <algorithm>
#include <complex>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#include <gnuradio/types.h>

namespace gr {
class io_signature
{
public:
    typedef std::shared_ptr<io_signature> sptr;
    int d_min, d_max;
    std::vector<int> d_sizes;
    static sptr make(int min_streams, int max_streams, int sizeof_stream_item)
    {
        sptr s(new io_signature);
        s->d_min = min_streams;
        s->d_max = max_streams;
        s->d_sizes.assign(std::max(max_streams, 1), sizeof_stream_item);
        return s;
    }
    static sptr makev(int min_streams, int max_streams, const std::vector<int>& sizes)
    {
        sptr s(new io_signature);
        s->d_min = min_streams;
        s->d_max = max_streams;
        s->d_sizes = sizes;
        return s;
    }
    int sizeof_stream_item(int i) const
    {
        if (d_sizes.empty()) return 0;
        return d_sizes[std::min<size_t>(i, d_sizes.size() - 1)];
    }
};
} // namespace gr
