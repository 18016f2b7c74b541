/* pybind11 bindings of the GPU-backed blocks: python classes ieee80211.signal(), ieee80211.signal2(),
 * ieee80211.demod(mupos, mugid), ieee80211.demod2() and ieee80211.decode(ifdebug), deriving gr::block / gr::basic_block exactly as the
 * reference's bindings do (python/ieee80211/bindings/{demod,demod2,decode}_python.cc:36-43), so
 * `from gnuradio import ieee80211` keeps working for rx.grc / rx2.grc.
 *
 * In the reference tree the functions below replace the bodies of bind_demod / bind_demod2 / bind_decode;
 * python_bindings.cc (which calls them) needs no change.  Not compiled in this repository: GNU Radio and
 * its pybind11 glue are not installed on the build or GPU boxes. */
#include <pybind11/pybind11.h>

#include <gnuradio/ieee80211/decode.h>
#include <gnuradio/ieee80211/detect.h>
#include <gnuradio/ieee80211/demod.h>
#include <gnuradio/ieee80211/demod2.h>
#include <gnuradio/ieee80211/signal.h>
#include <gnuradio/ieee80211/signal2.h>

namespace py = pybind11;

void bind_signal(py::module& m)
{
    using blk = ::gr::ieee80211::signal;
    py::class_<blk, gr::block, gr::basic_block, std::shared_ptr<blk>>(m, "signal", "802.11 legacy signal field and CFO correction (B200)")
        .def(py::init(&blk::make));
}

void bind_signal2(py::module& m)
{
    using blk = ::gr::ieee80211::signal2;
    py::class_<blk, gr::block, gr::basic_block, std::shared_ptr<blk>>(m, "signal2", "802.11 legacy signal field and CFO correction, two antennas (B200)")
        .def(py::init(&blk::make));
}

void bind_demod(py::module& m)
{
    using blk = ::gr::ieee80211::demod;
    py::class_<blk, gr::block, gr::basic_block, std::shared_ptr<blk>>(m, "demod", "802.11 OFDM demodulator (B200)")
        .def(py::init(&blk::make), py::arg("mupos"), py::arg("mugid"));
}

void bind_demod2(py::module& m)
{
    using blk = ::gr::ieee80211::demod2;
    py::class_<blk, gr::block, gr::basic_block, std::shared_ptr<blk>>(m, "demod2", "802.11 two-antenna OFDM demodulator (B200)")
        .def(py::init(&blk::make));
}

void bind_decode(py::module& m)
{
    using blk = ::gr::ieee80211::decode;
    py::class_<blk, gr::block, gr::basic_block, std::shared_ptr<blk>>(m, "decode", "802.11 BCC decoder (B200)")
        .def(py::init(&blk::make), py::arg("ifdebug"));
}

/* new block (no counterpart in the reference's bindings): add `void bind_detect(py::module&);` and a call to it in
 * python_bindings.cc when gr/lib/detect_impl.cc is built in */
void bind_detect(py::module& m)
{
    using blk = ::gr::ieee80211::detect;
    py::class_<blk, gr::block, gr::basic_block, std::shared_ptr<blk>>(m, "detect", "802.11 packet detection: presiso + trigger + sync in one block (B200)")
        .def(py::init(&blk::make));
}
