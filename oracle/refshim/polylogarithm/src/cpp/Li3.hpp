// oracle stand-in, see ../../polylog_neg.hpp
#include "../../polylog_neg.hpp"
