/* Mock of gr::block: exactly the surface the reference's demod / demod2 /
 * decode / signal / signal2 blocks use, plus the
 * hooks a single-threaded mini scheduler needs (mock_* members).
 * Test infrastructure only; it also hosts this repo's own drop-in blocks in
 * the block-level equivalence tests. */
#pragma once
#include <algorithm>
#include <chrono>
#include <iostream>
#include <cstdint>
#include <memory>
#include <mutex>
#include <string>
#include <utility>
#include <vector>
#include <gnuradio/io_signature.h>
#include <gnuradio/tags.h>
#include <pmt/pmt.h>

namespace gr {

/* One stream (an edge of the flowgraph): items + tags with absolute offsets. */
struct mock_stream {
    int itemsize;
    std::vector<uint8_t> data;
    std::vector<tag_t> tags;
    bool tags_sorted; /* tags are in ascending offset order: get_tags_in_range() may use a binary search */
    std::mutex* lock; /* set while two scheduler threads share the stream (mock_run_chain_threads): guards `tags` */
    mock_stream() : itemsize(1), tags_sorted(false), lock(nullptr) {}
    explicit mock_stream(int isz) : itemsize(isz), tags_sorted(false), lock(nullptr) {}
    uint64_t nitems() const { return itemsize ? data.size() / itemsize : 0; }
};

class basic_block
{
public:
    virtual ~basic_block() {}
};

class block : public basic_block
{
public:
    enum tag_propagation_policy_t { TPP_DONT = 0, TPP_ALL_TO_ALL = 1, TPP_ONE_TO_ONE = 2, TPP_CUSTOM = 3 };
    typedef std::shared_ptr<block> sptr;

    /* ---- mock scheduler state (public on purpose) ---- */
    std::string mock_name;
    io_signature::sptr mock_insig, mock_outsig;
    std::vector<const mock_stream*> mock_in; /* bound input streams */
    std::vector<mock_stream*> mock_out;      /* bound output streams */
    std::vector<uint64_t> mock_nread, mock_nwritten;
    std::vector<int> mock_consumed; /* per input, set by consume()/consume_each() in the current call */
    std::vector<std::pair<std::string, pmt::pmt_t>> mock_msgs; /* (port, message) in publication order */
    int mock_tpp;

protected:
    block() : mock_tpp(TPP_ALL_TO_ALL) {}
    block(const std::string& name, io_signature::sptr in, io_signature::sptr out)
        : mock_name(name), mock_insig(in), mock_outsig(out), mock_tpp(TPP_ALL_TO_ALL)
    {
    }

public:
    virtual ~block() {}
    virtual void forecast(int noutput_items, gr_vector_int& ninput_items_required)
    {
        for (auto& n : ninput_items_required) n = noutput_items;
    }
    virtual int general_work(int noutput_items,
                             gr_vector_int& ninput_items,
                             gr_vector_const_void_star& input_items,
                             gr_vector_void_star& output_items) = 0;
    virtual bool start() { return true; }
    virtual bool stop() { return true; }

    const std::string& name() const { return mock_name; }
    io_signature::sptr input_signature() const { return mock_insig; }
    io_signature::sptr output_signature() const { return mock_outsig; }

    void consume(int which_input, int how_many_items)
    {
        if ((size_t)which_input < mock_consumed.size()) mock_consumed[which_input] += how_many_items;
    }
    void consume_each(int how_many_items)
    {
        for (auto& c : mock_consumed) c += how_many_items;
    }
    uint64_t nitems_read(unsigned int which_input) { return mock_nread.at(which_input); }
    uint64_t nitems_written(unsigned int which_output) { return mock_nwritten.at(which_output); }

    void get_tags_in_range(std::vector<tag_t>& v, unsigned int which_input, uint64_t abs_start, uint64_t abs_end)
    {
        v.clear();
        const mock_stream* s = mock_in.at(which_input);
        std::unique_lock<std::mutex> guard;
        if (s->lock) guard = std::unique_lock<std::mutex>(*s->lock);
        if (s->tags_sorted) {
            auto it = std::lower_bound(s->tags.begin(), s->tags.end(), abs_start, [](const tag_t& t, uint64_t o) { return t.offset < o; });
            for (; it != s->tags.end() && it->offset < abs_end; ++it) v.push_back(*it);
            return;
        }
        for (const auto& t : s->tags)
            if (t.offset >= abs_start && t.offset < abs_end) v.push_back(t);
    }
    void add_item_tag(unsigned int which_output,
                      uint64_t abs_offset,
                      const pmt::pmt_t& key,
                      const pmt::pmt_t& value,
                      const pmt::pmt_t& srcid = pmt::PMT_NIL)
    {
        tag_t t;
        t.offset = abs_offset;
        t.key = key;
        t.value = value;
        t.srcid = srcid;
        mock_stream* o = mock_out.at(which_output);
        std::unique_lock<std::mutex> guard;
        if (o->lock) guard = std::unique_lock<std::mutex>(*o->lock);
        o->tags.push_back(t);
    }
    pmt::pmt_t alias_pmt() const { return pmt::mp(mock_name); }
    void set_tag_propagation_policy(tag_propagation_policy_t p) { mock_tpp = p; }
    tag_propagation_policy_t tag_propagation_policy() { return (tag_propagation_policy_t)mock_tpp; }
    void message_port_register_out(pmt::pmt_t /*port_id*/) {}
    void message_port_register_in(pmt::pmt_t /*port_id*/) {}
    void message_port_pub(pmt::pmt_t port_id, pmt::pmt_t msg)
    {
        mock_msgs.emplace_back(pmt::symbol_to_string(port_id), msg);
    }
    void set_output_multiple(int) {}
    void set_min_output_buffer(long) {}
};

} // namespace gr

namespace gnuradio {
template <class T, class... Args>
std::shared_ptr<T> make_block_sptr(Args&&... args)
{
    return std::shared_ptr<T>(new T(std::forward<Args>(args)...));
}
} // namespace gnuradio
