// message_classes.h — the six message classes of the control/data protocol.
// ref: /root/reference/src/core/system/message_classes.h:13-42 (same numbering).
#pragma once
namespace ssb {
enum MSG_CLS {
  NODE_INIT_ADDRESS = 0,         // node -> master: my address; reply (deferred): the route table
  NODE_ASKFOR_HASHFRAG = 1,      // node -> master: send me the fragment map
  WORKER_PULL_REQUEST = 2,       // worker -> server: keys; reply: (key, value)*
  WORKER_PUSH_REQUEST = 3,       // worker -> server: (key, grad)*; reply: ack
  WORKER_FINISH_WORK = 4,        // worker -> master: done
  SERVER_TOLD_TO_TERMINATE = 5   // master -> server: dump and exit
};
}  // namespace ssb
