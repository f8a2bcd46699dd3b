// rules.cu — the built-in update rules as named entry points (ssb_rule_<name>_info / _push / _push_bf16_seg).
// This file is what a user's own rule file looks like: the public header and one line per rule. FTRL, the third
// optimizer, exists only here and in the header — the table kernels (table.cu) know nothing about it.
// ref: core/parameter/sparse_access_method.h:21-46 (PullAccessMethod / PushAccessMethod are the user's classes).
#include "../../include/ssb200_rule.cuh"

SSB200_EXPORT_RULE(sgd, ssb::rule::Sgd)
SSB200_EXPORT_RULE(adagrad, ssb::rule::AdaGrad)
SSB200_EXPORT_RULE(ftrl, ssb::rule::Ftrl)
