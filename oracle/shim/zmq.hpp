/*
 * oracle/shim/zmq.hpp -- stand-in for cppzmq's <zmq.hpp> (not vendored by the reference).
 * TEST INFRASTRUCTURE ONLY.  A socket_t that records the last message it was asked to send, so
 * the harness can read the exact bytes cpacketize::send() publishes (src/cpacketizer.cc:109-129
 * of the reference tree) without any network.  Pulls in <vector>/<memory>, which the reference's
 * cpacketizer.h relies on transitively.
 */
#ifndef ORACLE_SHIM_ZMQ_HPP
#define ORACLE_SHIM_ZMQ_HPP
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#define ZMQ_PUB 1
#define ZMQ_SUB 2
namespace zmq {
class context_t {
public:
    explicit context_t(int = 1) {}
};
class socket_t {
public:
    std::vector<uint8_t> last;      // bytes of the most recent send()
    uint64_t n_sent = 0;
    std::string bound;
    socket_t(context_t &, int) {}
    void bind(const char *addr) { bound = addr; }
    void bind(const std::string &addr) { bound = addr; }
    size_t send(const void *buf, size_t len, int = 0) {
        last.assign((const uint8_t *)buf, (const uint8_t *)buf + len);
        n_sent++;
        return len;
    }
    void close() {}
};
}  // namespace zmq
#endif
