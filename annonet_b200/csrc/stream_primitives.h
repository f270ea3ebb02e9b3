// dlib stream primitives (dlib/dlib/serialize.h:297-345,130-131,598-632,1209-1221; float_details.h:86-135;
// cuda/tensor.h:444-466,657-665).  Internal to net_model/net_stream.
#pragma once
#include "net_model.h"
#include <cmath>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

namespace ab {

struct Reader {
    const unsigned char* p;
    size_t n, pos = 0;
    Reader(const void* b, size_t n_) : p((const unsigned char*)b), n(n_) {}

    void need(size_t k) const {
        if (k > n - pos) throw serialization_error("unexpected end of the serialized net at byte " + std::to_string(pos));   // (pos <= n always: no overflow)
    }
    long long integer() {  // serialize.h:297-345 (+ unpack_int :347-420)
        need(1);
        const unsigned c = p[pos];
        const unsigned len = c & 0x0F;
        if (len == 0 || len > 8) throw serialization_error("bad integer control byte at " + std::to_string(pos));
        need(1 + len);
        unsigned long long v = 0;
        for (unsigned i = 0; i < len; ++i) v |= (unsigned long long)p[pos + 1 + i] << (8 * i);
        pos += 1 + len;
        return (c & 0x80) ? -(long long)v : (long long)v;
    }
    bool boolean() {  // serialize.h:130-131
        need(1);
        const unsigned char c = p[pos++];
        if (c == '1') return true;
        if (c == '0') return false;
        throw serialization_error("bad bool byte at " + std::to_string(pos - 1));
    }
    double real() {  // float_details.h:86-135
        const long long m = integer();
        const long long e = integer();
        if (e == 32000) return INFINITY;
        if (e == 32001) return -INFINITY;
        if (e == 32002) return NAN;
        return std::ldexp((double)m, (int)e);
    }
    std::string string() {  // serialize.h:1209-1221
        const long long len = integer();
        if (len < 0) throw serialization_error("negative string length");
        need((size_t)len);
        std::string s((const char*)p + pos, (size_t)len);
        pos += (size_t)len;
        return s;
    }
    void tensor(HostTensor& t) {  // cuda/tensor.h:444-466
        if (integer() != 2) throw serialization_error("unexpected tensor version at " + std::to_string(pos));
        for (int i = 0; i < 4; ++i) t.d[i] = integer();
        // every dimension is checked before the product: a hostile stream must not overflow it (or pass a negative pair)
        unsigned long long cnt = 1;
        for (int i = 0; i < 4; ++i) {
            if (t.d[i] < 0) throw serialization_error("negative tensor dimension at " + std::to_string(pos));
            if (t.d[i] != 0 && cnt > (unsigned long long)(n / 4) / (unsigned long long)t.d[i])
                throw serialization_error("tensor larger than the serialized net at " + std::to_string(pos));
            cnt *= (unsigned long long)t.d[i];
        }
        need((size_t)cnt * 4);
        t.v.resize((size_t)cnt);
        if (cnt) std::memcpy(t.v.data(), p + pos, (size_t)cnt * 4);
        pos += (size_t)cnt * 4;
    }
    void alias(long long d[4]) {  // cuda/tensor.h:657-665
        if (integer() != 1) throw serialization_error("unexpected alias_tensor version at " + std::to_string(pos));
        for (int i = 0; i < 4; ++i) d[i] = integer();
    }
};

struct Writer {
    std::vector<char> out;
    void integer(long long v) {
        unsigned char buf[9];
        unsigned char neg = 0;
        unsigned long long u;
        if (v < 0) { neg = 0x80; u = (unsigned long long)(-v); } else u = (unsigned long long)v;
        unsigned char size = 8;
        for (unsigned char i = 1; i <= 8; ++i) {
            buf[i] = (unsigned char)(u & 0xFF);
            u >>= 8;
            if (u == 0) { size = i; break; }
        }
        buf[0] = size | neg;
        out.insert(out.end(), (char*)buf, (char*)buf + size + 1);
    }
    void boolean(bool b) { out.push_back(b ? '1' : '0'); }
    template <typename T>
    void real(T val) {
        long long mant = 0;
        int expo = 0;
        const int digits = std::numeric_limits<T>::digits;
        if (val == std::numeric_limits<T>::infinity()) expo = 32000;
        else if (val == -std::numeric_limits<T>::infinity()) expo = 32001;
        else if (val < std::numeric_limits<T>::infinity()) {
            int e;
            mant = (long long)(std::frexp(val, &e) * (double)(1ULL << digits));
            expo = e - digits;
            for (int i = 0; i < 8 && ((mant & 0xFF) == 0); ++i) { mant >>= 8; expo += 8; }
        } else expo = 32002;
        integer(mant);
        integer(expo);
    }
    void string(const std::string& s) {
        integer((long long)s.size());
        out.insert(out.end(), s.begin(), s.end());
    }
    void tensor(const HostTensor& t) {
        integer(2);
        for (int i = 0; i < 4; ++i) integer(t.d[i]);
        const char* b = (const char*)t.v.data();
        out.insert(out.end(), b, b + t.v.size() * 4);
    }
    void empty_tensor() { integer(2); for (int i = 0; i < 4; ++i) integer(0); }
    void alias(const long long d[4]) { integer(1); for (int i = 0; i < 4; ++i) integer(d[i]); }
};

}  // namespace ab
