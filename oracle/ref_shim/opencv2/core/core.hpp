// header shim, test infrastructure only: see ../opencv.hpp
#include "../opencv.hpp"
