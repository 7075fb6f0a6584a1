// CPU test of lo::CopyPool (csrc/lo_copy_pool.h): copies of many sizes, back to back (helpers spinning) and after pauses
// longer than the spin time (helpers asleep), checked byte for byte; construction / destruction with 0..4 helpers.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <thread>
#include <vector>

#include "../../lidar_odometry_demo_b200/csrc/lo_copy_pool.h"

int main() {
    std::mt19937_64 rng(7);
    std::vector<unsigned char> src(3u << 20), dst(3u << 20);
    for (auto& b : src) b = static_cast<unsigned char>(rng());
    long copies = 0;
    for (int helpers = 0; helpers <= 4; helpers++) {
        lo::CopyPool pool(helpers);
        for (int round = 0; round < 600; round++) {
            const size_t bytes = round % 7 == 0 ? 1 + rng() % 5000 : 1 + rng() % src.size();
            const size_t off = rng() % (src.size() - bytes + 1);
            std::fill(dst.begin() + off, dst.begin() + off + bytes, 0);
            if (round % 3 == 0) {
                pool.copy(dst.data() + off, src.data() + off, bytes);
            } else {   // the way lo_odometry_process uses it: consume the completed prefix piece by piece, help in between
                pool.begin(dst.data() + off, src.data() + off, bytes);
                size_t sent = 0;
                while (sent < bytes) {
                    const size_t ready = pool.ready_bytes();
                    if (ready < sent || ready > bytes) return 3;
                    if (ready - sent >= 256 * 1024 || (ready == bytes && ready > sent)) {
                        if (std::memcmp(dst.data() + off + sent, src.data() + off + sent, ready - sent) != 0) {
                            std::printf("prefix mismatch: helpers %d round %d bytes %zu at %zu..%zu\n", helpers, round, bytes, sent, ready);
                            return 4;
                        }
                        sent = ready;
                    } else if (!pool.help_one()) {
                        lo::CopyPool::cpu_relax();
                    }
                }
            }
            if (std::memcmp(dst.data() + off, src.data() + off, bytes) != 0) {
                std::printf("mismatch: helpers %d round %d bytes %zu\n", helpers, round, bytes);
                return 1;
            }
            if (off > 0 && dst[off - 1] != 0 && dst[off - 1] != src[off - 1]) return 2;
            copies++;
            if (round % 75 == 74) std::this_thread::sleep_for(std::chrono::microseconds(3 * lo::CopyPool::kSpinMicros));   // let the helpers fall asleep
        }
    }
    std::printf("ok %ld copies\n", copies);
    return 0;
}
