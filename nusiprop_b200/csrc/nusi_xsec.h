// nusi_xsec.h -- umbrella for the cross-section builder plus the table-row energy rule.
#ifndef NUSI_XSEC_H
#define NUSI_XSEC_H
#include "nusi_gamma.h"
#include "nusi_alphatilde.h"
#include "nusi_alpha.h"
#endif
