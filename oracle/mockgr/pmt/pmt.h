/* Mock of the subset of GNU Radio's PMT library that the reference blocks
 * demod / demod2 / decode / signal / signal2 touch.
 * Test infrastructure only: lets /root/reference/lib/*_impl.cc compile UNCHANGED
 * without GNU Radio.  Behaviour mirrored from the public PMT API:
 *   - dicts are association lists, dict_add puts the new pair in front and
 *     removes an older pair with the same key, dict_items returns that list;
 *   - symbols compare by string value;
 *   - from_float stores a double, to_float narrows it back.
 */
#pragma once
#include <complex>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace pmt {

enum kind_t { K_NIL, K_SYMBOL, K_LONG, K_DOUBLE, K_PAIR, K_C32VEC, K_BLOB };

struct pmt_base;
typedef std::shared_ptr<pmt_base> pmt_t;

struct pmt_base {
    kind_t kind;
    std::string sym;
    long l;
    double d;
    pmt_t car, cdr;
    std::vector<std::complex<float>> c32;
    std::vector<uint8_t> blob;
    explicit pmt_base(kind_t k) : kind(k), l(0), d(0.0) {}
};

inline const pmt_t PMT_NIL = std::make_shared<pmt_base>(K_NIL);

inline bool is_null(const pmt_t& p) { return !p || p->kind == K_NIL; }

inline pmt_t mp(const std::string& s)
{
    pmt_t p = std::make_shared<pmt_base>(K_SYMBOL);
    p->sym = s;
    return p;
}
inline pmt_t string_to_symbol(const std::string& s) { return mp(s); }
inline pmt_t intern(const std::string& s) { return mp(s); }
inline std::string symbol_to_string(const pmt_t& p)
{
    if (!p || p->kind != K_SYMBOL) throw std::runtime_error("pmt: not a symbol");
    return p->sym;
}

inline pmt_t from_long(long v)
{
    pmt_t p = std::make_shared<pmt_base>(K_LONG);
    p->l = v;
    return p;
}
inline long to_long(const pmt_t& p)
{
    if (p && p->kind == K_LONG) return p->l;
    if (p && p->kind == K_DOUBLE) return (long)p->d;
    throw std::runtime_error("pmt: to_long wrong type");
}
inline pmt_t from_double(double v)
{
    pmt_t p = std::make_shared<pmt_base>(K_DOUBLE);
    p->d = v;
    return p;
}
inline pmt_t from_float(float v) { return from_double((double)v); }
inline double to_double(const pmt_t& p)
{
    if (p && p->kind == K_DOUBLE) return p->d;
    if (p && p->kind == K_LONG) return (double)p->l;
    throw std::runtime_error("pmt: to_double wrong type");
}
inline float to_float(const pmt_t& p) { return (float)to_double(p); }

inline pmt_t cons(const pmt_t& a, const pmt_t& b)
{
    pmt_t p = std::make_shared<pmt_base>(K_PAIR);
    p->car = a;
    p->cdr = b;
    return p;
}
inline pmt_t car(const pmt_t& p)
{
    if (!p || p->kind != K_PAIR) throw std::runtime_error("pmt: car of non-pair");
    return p->car;
}
inline pmt_t cdr(const pmt_t& p)
{
    if (!p || p->kind != K_PAIR) throw std::runtime_error("pmt: cdr of non-pair");
    return p->cdr;
}
inline bool eqv(const pmt_t& a, const pmt_t& b)
{
    if (a == b) return true;
    if (!a || !b || a->kind != b->kind) return false;
    switch (a->kind) {
    case K_NIL: return true;
    case K_SYMBOL: return a->sym == b->sym;
    case K_LONG: return a->l == b->l;
    case K_DOUBLE: return a->d == b->d;
    default: return false;
    }
}

inline size_t length(const pmt_t& p)
{
    size_t n = 0;
    pmt_t q = p;
    while (q && q->kind == K_PAIR) {
        n++;
        q = q->cdr;
    }
    return n;
}
inline pmt_t nth(size_t n, const pmt_t& p)
{
    pmt_t q = p;
    while (n > 0 && q && q->kind == K_PAIR) {
        q = q->cdr;
        n--;
    }
    if (!q || q->kind != K_PAIR) return PMT_NIL;
    return q->car;
}

/* dictionaries as alists */
inline pmt_t make_dict() { return PMT_NIL; }
inline pmt_t dict_delete(const pmt_t& dict, const pmt_t& key)
{
    if (is_null(dict)) return PMT_NIL;
    if (eqv(car(car(dict)), key)) return cdr(dict);
    return cons(car(dict), dict_delete(cdr(dict), key));
}
inline bool dict_has_key(const pmt_t& dict, const pmt_t& key)
{
    pmt_t q = dict;
    while (q && q->kind == K_PAIR) {
        if (eqv(car(q->car), key)) return true;
        q = q->cdr;
    }
    return false;
}
inline pmt_t dict_add(const pmt_t& dict, const pmt_t& key, const pmt_t& value)
{
    if (is_null(dict)) return cons(cons(key, value), PMT_NIL);
    if (dict_has_key(dict, key)) return cons(cons(key, value), dict_delete(dict, key));
    return cons(cons(key, value), dict);
}
inline pmt_t dict_ref(const pmt_t& dict, const pmt_t& key, const pmt_t& not_found)
{
    pmt_t q = dict;
    while (q && q->kind == K_PAIR) {
        if (eqv(car(q->car), key)) return cdr(q->car);
        q = q->cdr;
    }
    return not_found;
}
inline pmt_t dict_items(const pmt_t& dict) { return dict; }

/* uniform vectors / blobs */
inline pmt_t init_c32vector(size_t k, const std::complex<float>* data)
{
    pmt_t p = std::make_shared<pmt_base>(K_C32VEC);
    p->c32.assign(data, data + k);
    return p;
}
inline pmt_t init_c32vector(size_t k, const std::vector<std::complex<float>>& data)
{
    return init_c32vector(k, data.data());
}
inline const std::vector<std::complex<float>> c32vector_elements(const pmt_t& p)
{
    /* real PMT throws on a non-vector; the reference passes PMT_NIL as the
     * default for a missing "chan" tag, which would throw there too. */
    if (!p || p->kind != K_C32VEC) throw std::runtime_error("pmt: c32vector_elements wrong type");
    return p->c32;
}
inline pmt_t make_blob(const void* buf, size_t len)
{
    pmt_t p = std::make_shared<pmt_base>(K_BLOB);
    p->blob.assign((const uint8_t*)buf, (const uint8_t*)buf + len);
    return p;
}
inline const void* blob_data(const pmt_t& p) { return p->blob.data(); }
inline size_t blob_length(const pmt_t& p) { return p->blob.size(); }

} // namespace pmt
