// graph_optimizer HOST PORT ITERATIONS PIPELINE SOLVER
//
// Same entry point as remote/app/main.cpp:5-47 (defaults 127.0.0.1 8888 10 cpu eigen; PIPELINE != "cpu"
// selects the GPU optimizer; PIPELINE=gpu forces SOLVER=cuda).  Transport: a dependency-free POSIX-socket
// loop reproducing remote/conn + remote/app/ConnectionHandlerGraph.h framing -- [int32 size][payload] both
// ways, connection kept open, one request at a time on one thread (main.cpp:42).  Boost.Asio is not in this
// image, and the transport is outside the accelerated path.
//
// Only the PIPELINE=gpu branch is shipped: `cpu eigen` is the reference's own Eigen pipeline, which this
// repository restates only as test infrastructure (the checker), never as product code.
#include <arpa/inet.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>

#include <chrono>
#include <csignal>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "graph_gpu.h"

namespace {

bool read_exact(int fd, void* buf, size_t n) {
    char* p = (char*)buf;
    while (n) { const ssize_t r = ::recv(fd, p, n, 0); if (r <= 0) return false; p += r; n -= (size_t)r; }
    return true;
}
bool write_all(int fd, const void* buf, size_t n) {
    const char* p = (const char*)buf;
    while (n) { const ssize_t r = ::send(fd, p, n, MSG_NOSIGNAL); if (r <= 0) return false; p += r; n -= (size_t)r; }
    return true;
}
struct Timer {   // remote/tools/BlockTimer.cpp:6-19
    std::string name; std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    explicit Timer(std::string n) : name(std::move(n)) {}
    ~Timer() { std::cout << name << ": " << std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() << " ms" << std::endl; }
};

// ConnectionHandler.h:14-34: deserialize -> optimize -> serialize -> send, then wait for the next size prefix
void serve(int fd, int iters) {
    std::cout << "\n------ New Connection ------\n";
    for (;;) {
        int32_t size = 0;
        if (!read_exact(fd, &size, 4)) return;                      // ConnectionHandlerGraph.h:16-22
        if (size <= 0) { std::cerr << "bad graph size " << size << "\n"; return; }
        std::vector<char> bytes((size_t)size);
        if (!read_exact(fd, bytes.data(), bytes.size())) return;    // :23-29
        Timer total("Total");
        try {
            OptimizerGpu<float> optimizer((unsigned)iters, nullptr);
            GraphGpu<float> graph;
            { Timer t("DeserializeGraph"); graph.LoadFromWire(bytes.data(), (uint32_t)bytes.size()); graph.ToDevice(GpuContext::Instance()); }
            { Timer t("OptimizeGPU"); optimizer.Optimize(&graph); }
            std::vector<uint8_t> data;
            { Timer t("SerializeGraph"); data = graph.Serialize(GpuContext::Instance()); }
            std::cout << "SendSync() data size: " << data.size() << std::endl;
            Timer t("Sending");
            if (!write_all(fd, data.data(), data.size())) return;
        } catch (const std::exception& e) {
            // the reference throws string literals nobody catches (DeserializeGraph.h:94-159); here a bad request
            // closes this connection and the server keeps running
            std::cerr << "request failed: " << e.what() << std::endl;
            return;
        }
    }
}

}  // namespace

int main(int argc, char* argv[]) {
    std::cout << "CUDA is supported\n";
    const std::string host = argc < 2 ? "127.0.0.1" : argv[1];
    const std::string port = argc < 3 ? "8888" : argv[2];
    const int iters = argc < 4 ? 10 : std::stoi(argv[3]);
    const std::string target = argc < 5 ? "cpu" : argv[4];
    const std::string solver = argc < 6 ? "eigen" : argv[5];
    const int optimizerType = target == "cpu" ? 0 : 1;
    int solverType = solver == "eigen" ? 0 : 1;
    if (optimizerType == 1) solverType = 1;                          // main.cpp:21-25
    std::cout << "iters: " << iters << ", optimizerType: " << optimizerType << ", solverType: " << solverType << std::endl;
    if (optimizerType == 0) {
        std::cerr << "this build ships the PIPELINE=gpu path only; `cpu eigen` is the reference's own pipeline\n";
        return 2;
    }
    try { const int n = GpuContext::Instance().world(); std::cout << "GPUs: " + std::to_string(n) << std::endl; } catch (const std::exception& e) { std::cerr << "ConnectionManager error: " << e.what() << std::endl; return 1; }

    addrinfo hints{}; hints.ai_family = AF_INET; hints.ai_socktype = SOCK_STREAM; hints.ai_flags = AI_PASSIVE;
    addrinfo* res = nullptr;
    if (getaddrinfo(host.c_str(), port.c_str(), &hints, &res) != 0 || !res) { std::cerr << "ConnectionManager error: resolve\n"; return 1; }
    const int ls = ::socket(res->ai_family, res->ai_socktype, res->ai_protocol);
    const int one = 1; setsockopt(ls, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    if (ls < 0 || ::bind(ls, res->ai_addr, res->ai_addrlen) != 0 || ::listen(ls, 4) != 0) {
        std::cerr << "ConnectionManager error: bind/listen " << std::strerror(errno) << std::endl; return 1; }
    freeaddrinfo(res);
    std::cout << "listening on " << host << ":" << port << std::endl;
    for (;;) {                                                        // ConnectionManagerServer.h:40-61
        const int fd = ::accept(ls, nullptr, nullptr);
        if (fd < 0) continue;
        setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
        serve(fd, iters);
        ::close(fd);
    }
}
