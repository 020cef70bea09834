// libvgio stand-in.  TEST INFRASTRUCTURE (see ../../README.md).  Graph loading is outside the path.
#pragma once
#include <memory>
#include <stdexcept>
#include <string>
namespace vg { namespace io {
struct VPKG {
    template <class T> static std::unique_ptr<T> load_one(const std::string&) {
        throw std::logic_error("oracle/_ref: libvgio stand-in called");
    }
};
}}  // namespace vg::io
