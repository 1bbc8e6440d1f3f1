// reference_usage.cpp -- TEST PROGRAM: every way the reference's own users call the one-call API, written against
// include/xsearch_b200/xsearch.hpp, so that "a user of the reference can switch" is a compiled and executed claim.
// The call shapes (not the code) come from the reference tree:
//   S1  README.md:33-41 (my_grep.cpp), example/grep.cpp:69-77   extern_search<lines>(pattern, file, ignore_case, num_threads), live loop
//   S2  checkit.cpp:3-7                                         extern_search<lines>(argv[1], argv[2], argv[3], 4, 2): char* meta path
//   S3  README.md:72                                            extern_search<count>(pattern, file, meta_file, num_threads, max_num_readers)
//   S4  test/src/xsearchTest.cpp:341-346                        extern_search<count_matches>(pattern, file, false, 1); join(); getResult()->size()
//   S5  test/src/xsearchTest.cpp:1226-1230                      extern_search<count_matches>(pattern, file, meta, ignore_case, 1, 1)
//   S6  test/src/xsearchTest.cpp:446-452                        join(); copyResultSafe(); std::sort on the copy
//   S7  test/src/xsearchTest.cpp:731-739                        live loop over a count: running totals, the last one is the total
//   S8  README.md:91                                            a searcher that is ignored: the destructor joins
// Built by tests/test_host_pipeline.py against the test double of the C ABI (and usable against libxsg.so on a GPU box).
// Prints one "Sk <value>" line per shape; the test compares them with the CPU oracle.
#include <xsearch_b200/xsearch.hpp>

#include <algorithm>
#include <iostream>
#include <numeric>
#include <string>

int main(int argc, char** argv) {
  if (argc != 4) return 64;
  const std::string pattern = argv[1], file_path = argv[2], meta_file_path = argv[3];
  const int num_threads = 2, max_num_readers = 2;
  bool ignore_case = false;

  {  // S1
    auto searcher = xs::extern_search<xs::lines>(argv[1], argv[2], false, 1);
    size_t n = 0, bytes = 0;
    for (auto const& line : *searcher->getResult()) {
      ++n;
      bytes += line.size();
    }
    std::cout << "S1 " << n << ' ' << bytes << '\n';
  }
  {  // S2: a char* third argument is the meta file, never the bool of the ignore_case overload
    auto t = xs::extern_search<xs::lines>(argv[1], argv[2], argv[3], 4, 2);
    size_t n = 0;
    for (auto i : *t->getResult()) n += 1 + 0 * i.size();
    std::cout << "S2 " << n << '\n';
  }
  {  // S3
    auto res = xs::extern_search<xs::count>(pattern, file_path, meta_file_path, num_threads, max_num_readers);
    res->join();
    std::cout << "S3 " << res->getResult()->size() << '\n';
  }
  {  // S4
    auto res = xs::extern_search<xs::count_matches>(pattern, file_path, false, 1);
    res->join();
    res->join();  // idempotent
    std::cout << "S4 " << res->getResult()->size() << '\n';
    auto icase = xs::extern_search<xs::count_lines>(pattern, file_path, true, 1);
    icase->join();
    std::cout << "S4i " << icase->getResult()->size() << '\n';
  }
  {  // S5
    ignore_case = true;
    auto res = xs::extern_search<xs::count_matches>(pattern, file_path, meta_file_path, ignore_case, 1, 1);
    res->join();
    std::cout << "S5 " << res->getResult()->size() << '\n';
  }
  {  // S6
    auto res = xs::extern_search<xs::line_byte_offsets>(pattern, file_path, false, 1);
    res->join();
    auto result = res->getResult()->copyResultSafe();
    const bool was_sorted = std::is_sorted(result.begin(), result.end());  // ours: always file order
    std::sort(result.begin(), result.end());
    std::cout << "S6 " << result.size() << ' ' << std::accumulate(result.begin(), result.end(), uint64_t{0}) << ' '
              << was_sorted << '\n';
    auto idx = xs::extern_search<xs::line_indices>(pattern, file_path, meta_file_path, false, 3, 3);
    idx->join();
    auto v = idx->getResult()->copyResultSafe();
    std::cout << "S6i " << v.size() << ' ' << std::accumulate(v.begin(), v.end(), uint64_t{0}) << '\n';
    auto off = xs::extern_search<xs::match_byte_offsets>(pattern, file_path);
    off->join();
    std::cout << "S6m " << off->getResult()->size() << '\n';
  }
  {  // S7
    auto res = xs::extern_search<xs::count_matches>(pattern, file_path, false, 1);
    uint64_t result = 0, steps = 0, monotone = 1;
    for (auto i : *res->getResult()) {
      monotone &= i >= result;
      result = i;
      ++steps;
    }
    std::cout << "S7 " << result << ' ' << (steps >= 1) << ' ' << monotone << '\n';
  }
  {  // S8
    xs::extern_search<xs::lines>(pattern, file_path, meta_file_path, num_threads, max_num_readers);
    auto kept = xs::extern_search<xs::count_lines>(pattern, file_path, meta_file_path, num_threads, max_num_readers);
    (void)kept->running();
  }
  std::cout << "S8 ok\n";
  return 0;
}
