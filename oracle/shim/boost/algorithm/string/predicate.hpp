// TEST SCAFFOLDING ONLY (oracle build). Minimal stand-in for the one Boost string
// predicate the reference includes (core/src/modules/include/IStructure.hpp:13,58).
// Boost is not installed in this image; this header is never part of the product library.
#pragma once
#include <cctype>
#include <string>
namespace boost {
namespace algorithm {
    inline bool iequals(const std::string& a, const std::string& b)
    {
        if (a.size() != b.size())
            return false;
        for (size_t k = 0; k < a.size(); ++k)
            if (std::tolower((unsigned char)a[k]) != std::tolower((unsigned char)b[k]))
                return false;
        return true;
    }
}
}
