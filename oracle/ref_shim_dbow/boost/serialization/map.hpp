// Header shim, TEST INFRASTRUCTURE ONLY: see serialization.hpp
#pragma once
#include "serialization.hpp"
