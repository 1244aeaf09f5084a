// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See se3.hpp.
#pragma once
#include "se3.hpp"
