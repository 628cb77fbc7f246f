// CPU ORACLE support (test infrastructure): stand-in for <boost/serialization/map.hpp>, see serialization.hpp beside it.
#pragma once
#include "serialization.hpp"
