// Minimal stand-in for <boost/shared_ptr.hpp> (Boost is absent from this image).  Test infrastructure only.
#pragma once
#include <memory>
namespace boost {
template <class T> using shared_ptr = std::shared_ptr<T>;
}
