// ORACLE / TEST INFRASTRUCTURE. The reference includes highgui (line2Dup.h:5) but the match path
// never calls it; nothing to provide (headless).
#ifndef SBM_ORACLE_SHIM_HIGHGUI_HPP
#define SBM_ORACLE_SHIM_HIGHGUI_HPP
#include "opencv2/core/core.hpp"
#endif
