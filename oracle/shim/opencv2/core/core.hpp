// ORACLE / TEST INFRASTRUCTURE. Stand-in for <opencv2/core/core.hpp> so that the reference's
// line2Dup.h:3 resolves; the container types come from the repo's own cv_compat header.
#ifndef SBM_ORACLE_SHIM_CORE_HPP
#define SBM_ORACLE_SHIM_CORE_HPP
#include "sbm/cv_compat.hpp"
#endif
