/* stand-in: see opencv.hpp */
#include "../opencv.hpp"
