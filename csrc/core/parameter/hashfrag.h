// hashfrag.h — hash-fragment key sharding: key -> fragment -> server node.
// ref: /root/reference/src/core/parameter/hashfrag.h:12-116 (BasicHashFrag).
// Same rule and wire form: fragment i belongs to node i / (frag_num / num_nodes) + 1 clamped
// to [1, num_nodes]; to_node_id(key) = map[fmix64(key) % frag_num]; serialised as
// int32 num_nodes, int32 num_frags, uint32 map[num_frags]. The indirection table is what
// would let fragments be re-homed; like the reference, nothing re-homes them.
// Differences by design: frag_num < num_nodes is rejected (the reference divides by zero,
// D11); one hashfrag instance serves every key type (the reference's singleton is keyed on
// index_t only, D2).
#pragma once
#include "../utils/all.h"
namespace ssb {

template <typename Key>
class BasicHashFrag {
 public:
  typedef Key key_t;
  BasicHashFrag() = default;

  // master side: frag_num from the config (ref: hashfrag.h:33), then assign fragments
  void init(int num_nodes) { init(num_nodes, global_config().get_config("frag_num").to_int32()); }
  void init(int num_nodes, int frag_num) {
    CHECK_GT(num_nodes, 0);
    CHECK_GE(frag_num, num_nodes) << "frag_num must be >= number of servers";
    num_nodes_ = num_nodes;
    num_frags_ = frag_num;
    map_.assign(frag_num, 0);
    int per = frag_num / num_nodes;
    for (int i = 0; i < frag_num; ++i) {
      int node = i / per + 1;
      if (node > num_nodes) node = num_nodes;
      map_[i] = (index_t)node;
    }
  }
  bool inited() const { return !map_.empty(); }
  int to_frag_id(const key_t& key) const { return (int)(get_hash_code((uint64_t)key) % (uint64_t)num_frags_); }
  int to_node_id(const key_t& key) const {
    CHECK(!map_.empty()) << "hash fragment map has not been initialised";
    return (int)map_[to_frag_id(key)];
  }
  int num_nodes() const { return num_nodes_; }
  int num_frags() const { return num_frags_; }
  const std::vector<index_t>& map_table() const { return map_; }

  void serialize(BinaryBuffer& bb) const {
    bb << (int32_t)num_nodes_ << (int32_t)num_frags_;
    bb.put_array(map_.data(), map_.size());
  }
  void deserialize(BinaryBuffer& bb) {
    int32_t n = 0, f = 0;
    bb >> n >> f;
    CHECK_GT(n, 0);
    CHECK_GT(f, 0);
    num_nodes_ = n;
    num_frags_ = f;
    map_.resize(f);
    bb.get_array(map_.data(), (size_t)f);
  }
  friend std::ostream& operator<<(std::ostream& os, const BasicHashFrag& h) {
    os << "hashfrag nodes=" << h.num_nodes_ << " frags=" << h.num_frags_;
    return os;
  }

 private:
  int num_nodes_ = 0, num_frags_ = 0;
  std::vector<index_t> map_;
};

template <typename Key>
inline BasicHashFrag<Key>& global_hashfrag() {
  static BasicHashFrag<Key> h;
  return h;
}
// the hash-fragment table is key-type independent: every instantiation shares one map
inline BasicHashFrag<uint64_t>& shared_hashfrag() { return global_hashfrag<uint64_t>(); }

}  // namespace ssb
