// Stand-in for boost/lexical_cast.hpp (Boost is not installed in this image).  Only used
// to compile the upstream program for the measured CUDA baseline (oracle/_ref); it covers
// the single use there: boost::lexical_cast<int>(char*) for the command-line integers.
#pragma once
#include <cstdlib>
#include <stdexcept>
namespace boost {
template <class T> inline T lexical_cast(const char* s) {
    char* end = nullptr;
    long v = std::strtol(s, &end, 10);
    if (end == s || *end != '\0') throw std::runtime_error("bad lexical cast");
    return (T)v;
}
}  // namespace boost
