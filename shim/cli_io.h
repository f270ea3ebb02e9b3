// TEST INFRASTRUCTURE ONLY (oracle). Raw little-endian file helpers of the `annonet_ref` driver.
#pragma once
#include <dlib/dnn.h>
#include <cstdint>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace ref_io {

// record = int32 layer index, int64 n,k,nr,nc, then n*k*nr*nc float32 (NCHW, dlib order)
inline void write_tensor(std::ostream& out, int idx, const dlib::tensor& t)
{
    const int32_t i = idx;
    const int64_t dims[4] = {t.num_samples(), t.k(), t.nr(), t.nc()};
    out.write(reinterpret_cast<const char*>(&i), sizeof(i));
    out.write(reinterpret_cast<const char*>(dims), sizeof(dims));
    if (t.size() != 0)
        out.write(reinterpret_cast<const char*>(t.host()), sizeof(float) * t.size());
}

inline std::vector<char> slurp(const std::string& path)
{
    std::ifstream in(path, std::ios::binary);
    if (!in) throw std::runtime_error("cannot open " + path);
    return std::vector<char>((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
}

}  // namespace ref_io
