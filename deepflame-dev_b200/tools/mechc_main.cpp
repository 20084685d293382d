// dfchem_mechc: command-line front end of the host mechanism compiler.
//   dfchem_mechc <mechanism.yaml> <out.blob>     compile to the flat DFCHEMB1 blob
//   dfchem_mechc <mechanism.yaml>                print a summary
#include <cstdio>
#include <exception>
#include <fstream>

#include "../csrc/mechanism.hpp"

int main(int argc, char** argv) {
    if (argc < 2) {
        std::fprintf(stderr, "usage: %s <mechanism.yaml> [out.blob]\n", argv[0]);
        return 2;
    }
    try {
        dfchem::Mechanism m = dfchem::compile_mechanism_file(argv[1]);
        if (argc >= 3) {
            std::vector<uint8_t> b = m.serialize();
            std::ofstream f(argv[2], std::ios::binary);
            f.write(reinterpret_cast<const char*>(b.data()), (std::streamsize)b.size());
            if (!f) {
                std::fprintf(stderr, "cannot write %s\n", argv[2]);
                return 1;
            }
        } else {
            int n[4] = {0, 0, 0, 0};
            for (int k : m.kind) ++n[k];
            std::printf("%d species, %d reactions (%d elementary, %d three-body, %d Lindemann, %d Troe)\n", m.N, m.R,
                        n[0], n[1], n[2], n[3]);
        }
    } catch (const std::exception& e) {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}
