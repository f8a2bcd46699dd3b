// message_classes.h — the six message classes of the control/data protocol (same numbering as the reference), plus
// one of this framework's own for GPU jobs (group_exchange.h).
// ref: /root/reference/src/core/system/message_classes.h:13-42.
#pragma once
namespace ssb {
enum MSG_CLS {
  NODE_INIT_ADDRESS = 0,         // node -> master: my address; reply (deferred): the route table
  NODE_ASKFOR_HASHFRAG = 1,      // node -> master: send me the fragment map
  WORKER_PULL_REQUEST = 2,       // worker -> server: keys; reply: (key, value)*
  WORKER_PUSH_REQUEST = 3,       // worker -> server: (key, grad)*; reply: ack
  WORKER_FINISH_WORK = 4,        // worker -> master: done
  SERVER_TOLD_TO_TERMINATE = 5,  // master -> server: dump and exit
  NODE_GROUP_EXCHANGE = 6        // GPU rank -> master: my bytes for round `seq`; reply (deferred): every rank's bytes
};
}  // namespace ssb
