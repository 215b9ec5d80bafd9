"""Regenerates the committed golden fixtures from the reference's own tests.

Run in the build container (needs /root/reference):  python tests/golden/make_golden.py
The fixtures are *data* the reference's tests hold for this path:
  snappy_app_{1,2}.comp / .raw   <- tests/test_snappy_app.cpp:209-222
                                    (comp_data1/2, decomp_data_expected1/2; only the
                                    first 497 / 434 compressed and 709 / 7581
                                    uncompressed bytes are meaningful, as in the test)
"""
import os
import re

REF = "/root/reference/tests/test_snappy_app.cpp"
HERE = os.path.dirname(os.path.abspath(__file__))


def c_array(text, name):
    m = re.search(r"uint8_t\s+" + name + r"\[\]\s*=\s*\{([^}]*)\}", text)
    return bytes(int(x, 16) for x in re.findall(r"0x[0-9A-Fa-f]+", m.group(1)))


def main():
    text = open(REF).read()
    sizes = {n: int(v) for n, v in re.findall(r"constexpr size_t (\w+) = (\d+);", text)}
    for i in (1, 2):
        comp = c_array(text, f"comp_data{i}")[: sizes[f"size_comp_data{i}"]]
        raw = c_array(text, f"decomp_data_expected{i}")[: sizes[f"uncomp_output_size{i}"]]
        open(os.path.join(HERE, f"snappy_app_{i}.comp"), "wb").write(comp)
        open(os.path.join(HERE, f"snappy_app_{i}.raw"), "wb").write(raw)
        print(i, len(comp), len(raw))


if __name__ == "__main__":
    main()
