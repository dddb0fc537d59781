// C++ counterpart of integration/go/pkg/xet/xetgpu_test.go: the host-side mirror (include/xetgpu.hpp) against
// hf_xet golden values.  Built as tools/xetgpu_host_test, run by tests/test_gpu_parity.py.
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <string>
#include <vector>

#include "../../include/xetgpu.hpp"

static int failures = 0;
#define EXPECT(cond)                                                   \
    do {                                                               \
        if (!(cond)) {                                                 \
            std::fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond); \
            failures++;                                                \
        }                                                              \
    } while (0)

int main(int argc, char** argv) {
    const std::string dir = argc > 1 ? argv[1] : "/tmp";
    struct Case {
        const char* name;
        std::string data;
        const char* hash;
    };
    std::vector<Case> cases = {
        {"empty", "", "0000000000000000000000000000000000000000000000000000000000000000"},
        {"hello_world", "hello world", "f15edd43b7e85cfd25e2758ffe13b99fdd09dfa817e8d9f6548722a7a126db7f"},
        {"zeros_8192", std::string(8192, '\0'), "711574865581cce65f5d06a1818a37a1dd4cfe3f65e3f4aaae2b1bacbfc253db"},
        {"zeros_131073", std::string(131073, '\0'), "83f8f48adc7310b5748295b256ca24cdce2aac457679c98526e3a19e0388f58a"},
        {"zeros_1MiB", std::string(1 << 20, '\0'), "1e671fe124cea35586b1d1c30b9d4fc6b4e05ee60c93406986444f7c23d54056"},
    };
    try {
        xet::GpuHasher h(-1);
        std::vector<std::string> paths, hashes;
        for (auto& c : cases) {
            std::string p = dir + "/" + c.name + ".bin";
            std::ofstream(p, std::ios::binary).write(c.data.data(), (std::streamsize)c.data.size());
            paths.push_back(p);
            hashes.push_back(c.hash);
            EXPECT(h.HashBytes(c.data.data(), c.data.size()).Hash == c.hash);
        }
        int progress_calls = 0;
        h.SetProgressHandler([&](const xet::ProgressUpdate& u) { progress_calls += u.TotalFiles == cases.size(); }, 0);
        auto digests = h.HashFiles(paths);
        h.SetProgressHandler(nullptr);
        EXPECT(progress_calls > 0);
        for (size_t i = 0; i < cases.size(); i++) {
            EXPECT(digests[i].Hash == cases[i].hash);
            EXPECT(digests[i].Size == cases[i].data.size());
            h.VerifyFile(paths[i], cases[i].hash);
        }
        EXPECT(digests[3].Chunks == 2);
        auto ok = h.VerifyFiles(paths, hashes);
        EXPECT(ok.size() == cases.size() && ok[0] && ok[4]);
        // a wrong hash is ChecksumMismatch (6) with both hashes in Details, as VerifyFile in xetgpu.go
        try {
            h.VerifyFile(paths[1], std::string(64, '0'));
            EXPECT(false);
        } catch (const xet::XetError& e) {
            EXPECT(xet::IsChecksumMismatch(e));
            EXPECT(e.Details.find(cases[1].hash) != std::string::npos);
        }
        hashes[2] = std::string(64, 'f');
        hashes[0] = "";  // skipped
        try {
            h.VerifyFiles(paths, hashes);
            EXPECT(false);
        } catch (const xet::XetError& e) {
            EXPECT(e.Code == 6);
            EXPECT(h.LastVerdicts()[0] && h.LastVerdicts()[1] && !h.LastVerdicts()[2]);
        }
        try {
            h.HashFiles({dir + "/does-not-exist"});
            EXPECT(false);
        } catch (const xet::XetError& e) {
            EXPECT(e.Code == XETGPU_ERR_NOT_FOUND);
        }
        try {
            h.HashFiles(paths, [] { return true; });
            EXPECT(false);
        } catch (const xet::XetError& e) {
            EXPECT(e.Code == XETGPU_ERR_CANCELLED);
        }
        try {
            h.VerifyFile(paths[0], "xyz");
            EXPECT(false);
        } catch (const xet::XetError& e) {
            EXPECT(e.Code == XETGPU_ERR_INVALID_CONFIG);
        }
        // the "xet" case of the replica agent's GetFileChecksum: path + algorithm in, digest string out
        EXPECT(xet::GetFileChecksum(h, paths[1], xet::XetChecksumAlgorithm) == cases[1].hash);
        EXPECT(h.XetFileChecksum(paths[4]) == cases[4].hash);
        EXPECT(std::string(xet::XetChecksumMetadataKey) == "opc-meta-xet");
        try {
            xet::GetFileChecksum(h, paths[1], "crc32");
            EXPECT(false);
        } catch (const std::invalid_argument& e) {
            EXPECT(std::string(e.what()) == "unsupported checksum algorithm: crc32");
        }
        try {
            h.XetFileChecksum(dir + "/does-not-exist");
            EXPECT(false);
        } catch (const xet::XetError& e) {
            EXPECT(e.Code == XETGPU_ERR_NOT_FOUND);
        }
        h.Close();
        // the node-wide hasher a single model-agent process would hold (all visible GPUs)
        xet::GpuMultiHasher m;
        EXPECT(m.Devices() >= 1);
        for (size_t i = 0; i < cases.size(); i++) hashes[i] = cases[i].hash;
        auto md = m.HashFiles(paths);
        for (size_t i = 0; i < cases.size(); i++) EXPECT(md[i].Hash == cases[i].hash);
        auto mok = m.VerifyFiles(paths, hashes);
        EXPECT(mok.size() == cases.size() && mok[0] && mok[3]);
        hashes[3] = std::string(64, '1');
        try {
            m.VerifyFiles(paths, hashes);
            EXPECT(false);
        } catch (const xet::XetError& e) {
            EXPECT(xet::IsChecksumMismatch(e) && !m.LastVerdicts()[3] && m.LastVerdicts()[4]);
        }
        m.Close();
    } catch (const xet::XetError& e) {
        std::fprintf(stderr, "unexpected: %s\n", e.what());
        return 2;
    }
    std::printf("%s (%d failures)\n", failures ? "FAILED" : "ok", failures);
    return failures ? 1 : 0;
}
