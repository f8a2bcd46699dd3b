// swiftsnails.h — umbrella header of the host core.
// ref: /root/reference/src/swiftsnails.h:7-17 (exports utils, param cache, pull/push access,
// sparse table, access methods and the three role classes).
#pragma once
#include "framework/master.h"
#include "framework/server.h"
#include "framework/worker.h"
#include "parameter/access_method.h"
#include "parameter/hashfrag.h"
#include "parameter/param_cache.h"
#include "parameter/pull_access.h"
#include "parameter/push_access.h"
#include "parameter/sparsetable.h"
#include "thread_pool.h"
#include "utils/all.h"
