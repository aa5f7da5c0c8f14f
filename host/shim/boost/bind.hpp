// Minimal stand-in for <boost/bind.hpp> (oracle build only).
#pragma once
#include <functional>
namespace boost { using std::bind; }
