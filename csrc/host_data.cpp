// Host-side runtime of libdtf_b200: dataset file readers.
//
// The reference decodes CIFAR-10 records inside tf.data (C++: FixedLengthRecordDataset + decode_raw +
// transpose, cifar10_main.py:54-109) and MNIST through numpy (mnist_model.py:128-150).  Here the whole
// dataset is read ONCE into a pinned / device-resident uint8 tensor (augmentation and standardisation
// happen on the GPU, per step), so the host work is: read the files, split label / pixels, turn the
// CHW planes of a CIFAR record into HWC.  That is done here with a few threads and no intermediate
// copies; the Python fallback in utils/data/datasets.py does the same with numpy.
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <thread>
#include <vector>

#include "dtf_b200.h"

namespace {
constexpr int kCifarRecord = 1 + 3 * 32 * 32;

void cifar_range(const uint8_t* rec, long long begin, long long end, uint8_t* images, int* labels) {
    for (long long i = begin; i < end; ++i) {
        const uint8_t* r = rec + i * kCifarRecord;
        labels[i] = r[0];
        const uint8_t* R = r + 1;
        const uint8_t* G = R + 1024;
        const uint8_t* B = G + 1024;
        uint8_t* o = images + i * 3072;
        for (int p = 0; p < 1024; ++p) {
            o[3 * p] = R[p];
            o[3 * p + 1] = G[p];
            o[3 * p + 2] = B[p];
        }
    }
}
}  // namespace

// records: n x 3073 bytes (label, R plane, G plane, B plane).  images: n x 32 x 32 x 3, labels: n int32.
extern "C" int dtf_cifar_decode(const uint8_t* records, long long n, uint8_t* images, int* labels, int threads) {
    if (records == nullptr || images == nullptr || labels == nullptr || n < 0) return DTF_ERR_ARG;
    if (threads < 1) threads = 1;
    if (threads > 32) threads = 32;
    if (n < 1024 || threads == 1) {
        cifar_range(records, 0, n, images, labels);
        return DTF_OK;
    }
    std::vector<std::thread> pool;
    const long long per = (n + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
        const long long b = t * per, e = b + per < n ? b + per : n;
        if (b >= e) break;
        pool.emplace_back(cifar_range, records, b, e, images, labels);
    }
    for (auto& th : pool) th.join();
    return DTF_OK;
}

// Reads exactly `nbytes` from `path` at `offset` into `dst`.  Returns the number of bytes read or a DTF_ERR_* code.
extern "C" long long dtf_read_file(const char* path, long long offset, void* dst, long long nbytes) {
    if (path == nullptr || dst == nullptr || nbytes < 0 || offset < 0) return DTF_ERR_ARG;
    FILE* f = fopen(path, "rb");
    if (f == nullptr) return DTF_ERR_ARG;
    if (offset > 0 && fseek(f, static_cast<long>(offset), SEEK_SET) != 0) {
        fclose(f);
        return DTF_ERR_ARG;
    }
    const size_t got = fread(dst, 1, static_cast<size_t>(nbytes), f);
    fclose(f);
    return static_cast<long long>(got);
}

// idx (MNIST) header: magic 0x0000 08 <ndim>, then ndim big-endian uint32 dims.  Fills dims[0..ndim) and returns
// the byte offset of the payload, or a DTF_ERR_* code.
extern "C" int dtf_idx_header(const uint8_t* buf, long long nbytes, int* ndim, int* dims, int max_dims) {
    if (buf == nullptr || ndim == nullptr || dims == nullptr || nbytes < 4) return DTF_ERR_ARG;
    if (buf[0] != 0 || buf[1] != 0 || buf[2] != 0x08) return DTF_ERR_SHAPE;     // unsigned byte payloads only
    const int nd = buf[3];
    if (nd < 1 || nd > max_dims || nbytes < 4 + 4LL * nd) return DTF_ERR_SHAPE;
    long long total = 1;
    for (int i = 0; i < nd; ++i) {
        const uint8_t* p = buf + 4 + 4 * i;
        dims[i] = (p[0] << 24) | (p[1] << 16) | (p[2] << 8) | p[3];
        total *= dims[i];
    }
    if (nbytes < 4 + 4LL * nd + total) return DTF_ERR_SHAPE;
    *ndim = nd;
    return 4 + 4 * nd;
}
