// Minimal stand-in for Dune::ParameterTreeParser (dune-common absent here):
// ini files ([section] / key = value / # comments) and "-key value" command
// line options, as used by fieldgenerator.cc:259-263 of the reference.
#ifndef PF_STANDIN_DUNE_COMMON_PARAMETERTREEPARSER_HH
#define PF_STANDIN_DUNE_COMMON_PARAMETERTREEPARSER_HH

#include <fstream>
#include <istream>
#include <set>
#include <string>

#include <dune/common/parametertree.hh>

namespace Dune {

class ParameterTreeParser
{
  static std::string ltrim(const std::string& s)
  {
    std::size_t i = 0;
    while (i < s.size() && (s[i] == ' ' || s[i] == '\t' || s[i] == '\n' ||
                            s[i] == '\r'))
      i++;
    return s.substr(i);
  }
  static std::string rtrim(const std::string& s)
  {
    std::size_t n = s.size();
    while (n > 0 && (s[n - 1] == ' ' || s[n - 1] == '\t' || s[n - 1] == '\n' ||
                     s[n - 1] == '\r'))
      n--;
    return s.substr(0, n);
  }

public:
  static void readINITree(std::istream& in, ParameterTree& pt,
                          bool overwrite = true)
  {
    readINITree(in, pt, "stream", overwrite);
  }

  static ParameterTree readINITree(const std::string& file)
  {
    ParameterTree pt;
    readINITree(file, pt, true);
    return pt;
  }

  static void readINITree(std::string file, ParameterTree& pt,
                          bool overwrite = true)
  {
    std::ifstream in(file);
    if (!in)
      DUNE_THROW(IOError, "Could not open configuration file " << file);
    readINITree(in, pt, "file '" + file + "'", overwrite);
  }

  static void readINITree(std::istream& in, ParameterTree& pt,
                          const std::string srcname, bool overwrite)
  {
    std::string prefix;
    std::set<std::string> keysInFile;
    while (!in.eof()) {
      std::string line;
      getline(in, line);
      line = ltrim(line);
      if (line.size() == 0)
        continue;
      switch (line[0]) {
        case '#':
          break;
        case '[': {
          line = rtrim(line);
          if (line[line.length() - 1] == ']') {
            prefix = rtrim(ltrim(line.substr(1, line.length() - 2)));
            if (prefix != "")
              prefix += ".";
          }
          break;
        }
        default: {
          const auto comment = line.find('#');
          line = line.substr(0, comment);
          const auto mid = line.find('=');
          if (mid == std::string::npos)
            break;
          const std::string key = prefix + rtrim(ltrim(line.substr(0, mid)));
          std::string value = ltrim(line.substr(mid + 1));
          if (value.length() > 0) {
            if (value[0] == '\'' || value[0] == '"') {
              const char quote = value[0];
              value = value.substr(1);
              while (*(rtrim(value).end() - 1) != quote) {
                if (!in.eof()) {
                  std::string l;
                  getline(in, l);
                  value = value + "\n" + l;
                } else
                  value = value + quote;
              }
              value = rtrim(value);
              value = value.substr(0, value.length() - 1);
            } else
              value = rtrim(value);
          }
          if (keysInFile.count(key) != 0)
            DUNE_THROW(ParameterTreeParserError,
                       "Key '" << key << "' appears twice in " << srcname
                               << " !");
          if (overwrite || !pt.hasKey(key))
            pt[key] = value;
          keysInFile.insert(key);
        }
      }
    }
  }

  static void readOptions(int argc, char* argv[], ParameterTree& pt)
  {
    for (int i = 1; i < argc; i++) {
      if ((argv[i][0] == '-') && (argv[i][1] != '\0')) {
        if (argv[i + 1] == nullptr || i + 1 >= argc)
          DUNE_THROW(RangeError, "last option on command line ("
                                   << argv[i]
                                   << ") does not have an argument");
        pt[argv[i] + 1] = argv[i + 1];
        ++i;
      }
    }
  }
};

} // namespace Dune
#endif
