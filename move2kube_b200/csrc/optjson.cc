// m2k_option_json: what an accepting detect script prints on stdout, from the arrays of a detection result.
//
// At translate time the reference runs the chosen option's detect script AGAIN on the service directory
// just to read this JSON (DockerfileContainerizer.GetContainer internal/containerizer/dockerfilecontainerizer.go:105-129,
// S2IContainerizer.GetContainer s2icontainerizer.go:97-115) -- for the python options that is one more
// `grep -R` over every byte under the directory.  Everything the 18 built-in scripts print is either a
// constant or one of two things the detection pass already holds: the first `*.war` entry of the
// directory (internal/assets/dockerfiles/java-war-*/m2kdfdetect.sh:20-26) and the main script found
// by the byte scan (dockerfiles/python/m2kdfdetect.sh:21-23, s2i/python/m2ks2idetect.sh:22-24).
// Bytes are reproduced exactly as the scripts write them (`echo` adds a newline, `printf` does not;
// names are inserted unescaped, as the scripts do).  Host code: no GPU involved.
#include <string.h>

#include <string>

#include "m2k_internal.h"

namespace {

// basename of the entry `"$1/"*.war` expands to first: non-dot names ending ".war", byte order (LC_ALL=C)
bool first_war(const m2k_tree* t, uint32_t d, std::string* name) {
  for (uint32_t e = t->dir_first_entry[d]; e < t->dir_first_entry[d + 1]; e++) {
    const uint32_t a = t->entry_name_off[e], b = t->entry_name_off[e + 1];
    if (b - a >= 4 && t->names[a] != '.' && memcmp(t->names + b - 4, ".war", 4) == 0) {
      name->assign(reinterpret_cast<const char*>(t->names + a), b - a);
      return true;
    }
  }
  return false;
}

// path of file f relative to directory d (f lies in d's subtree): what `realpath --relative-to` prints
// when no symbolic link is involved
std::string rel_path(const m2k_tree* t, uint32_t f, uint32_t d) {
  const uint32_t e = t->file_entry[f];
  std::string out(reinterpret_cast<const char*>(t->names + t->entry_name_off[e]), t->entry_name_off[e + 1] - t->entry_name_off[e]);
  for (uint32_t x = t->entry_dir[e]; x != d && x != M2K_NONE && x != 0; x = t->dir_parent[x]) {
    const uint32_t de = t->dir_entry[x];
    out.insert(0, "/");
    out.insert(0, reinterpret_cast<const char*>(t->names + t->entry_name_off[de]), t->entry_name_off[de + 1] - t->entry_name_off[de]);
  }
  return out;
}

}  // namespace

extern "C" int m2k_option_json(const m2k_tree* tree, const m2k_result* res, uint32_t dir, uint32_t option, char* buf, size_t cap) {
  if (!tree || !res || dir >= tree->n_dirs || dir >= res->n_dirs || option >= M2K_N_OPTS) {
    m2k::set_global_error("m2k_option_json: bad argument");
    return M2K_ERR_ARG;
  }
  if (!((res->dir_opts[dir] >> option) & 1u)) {
    m2k::set_global_error("m2k_option_json: the option's detect script does not accept this directory (it prints nothing and exits 1)");
    return M2K_ERR_ARG;
  }
  std::string s;
  auto main_script = [&]() {
    const uint32_t f = res->dir_main_file[dir];
    return f == M2K_NONE ? std::string() : rel_path(tree, f, dir);
  };
  switch (option) {
    case M2K_DF_DJANGO: s = "{\"port\": 8080, \"binding\": \"0.0.0.0:8080\"}\n"; break;
    case M2K_DF_GOLANG: s = "{\"port\": 8080, \"app_name\": \"app-bin\"}\n"; break;
    case M2K_DF_WAR_JBOSS: case M2K_DF_WAR_LIBERTY: case M2K_DF_WAR_TOMCAT: {
      std::string war;
      if (!first_war(tree, dir, &war)) { m2k::set_global_error("m2k_option_json: no *.war entry in the directory"); return M2K_ERR_ARG; }
      s = std::string("{\"port\":") + (option == M2K_DF_WAR_LIBERTY ? "9080" : "8080") + ", \"war_path\":\"" + war + "\"}";
      break;
    }
    case M2K_DF_JAVAANT: s = "{\"port\": 8080, \"ant_cmd\": \"ant all\", \"app_name\": \"simplewebapp\"}\n"; break;
    case M2K_DF_JAVAGRADLE: s = "{\"port\": 8080, \"app_name\": \"simplewebapp\"}\n"; break;
    case M2K_DF_JAVAMAVEN: case M2K_DF_NODEJS: case M2K_DF_RUBY: s = "{\"port\": 8080, \"app_name\": \"app\"}\n"; break;
    case M2K_DF_PHP: s = "{\"port\": 8080, \"binding\": \"0.0.0.0:8080\", \"app_name\": \"app\"}\n"; break;
    case M2K_DF_PYTHON: s = "{\"main_script_rel_path\": \"" + main_script() + "\", \"app_name\": \"app\", \"port\": 8080}"; break;
    case M2K_S2I_GOLANG: s = "{\"builder\": \"registry.access.redhat.com/ubi8/go-toolset:latest\", \"port\": 8080}\n"; break;
    case M2K_S2I_JAVA:
      // pom.xml -> the web image, else (a *.java somewhere below) the native image: s2i/java/m2ks2idetect.sh:31-41
      s = std::string("{\"builder\": \"registry.access.redhat.com/") +
          (((res->dir_direct[dir] >> M2K_F_POM_XML) & 1u) ? "jboss-eap-6/eap64-openshift:latest" : "redhat-openjdk-18/openjdk18-openshift:latest") +
          "\", \"port\": 8080}\n";
      break;
    case M2K_S2I_NODEJS: s = "{\"builder\": \"registry.access.redhat.com/ubi8/nodejs-10\", \"port\": 8080}\n"; break;
    case M2K_S2I_PHP: s = "{\"builder\": \"registry.access.redhat.com/rhscl/php-72-rhel7:latest\", \"port\": 8080}\n"; break;
    case M2K_S2I_PYTHON:
      s = "{\"builder\": \"registry.access.redhat.com/rhscl/python-36-rhel7:latest\", \"app_file\": \"" + main_script() +
          "\", \"app_name\": \"app\", \"port\": 8080}";
      break;
    case M2K_S2I_RUBY: s = "{\"builder\": \"registry.access.redhat.com/rhscl/ruby-25-rhel7:latest\", \"port\": 8080}\n"; break;
    default: return M2K_ERR_ARG;
  }
  if (buf && cap) {
    if (s.size() + 1 > cap) { m2k::set_global_error("m2k_option_json: buffer too small"); return M2K_ERR_LIMIT; }
    memcpy(buf, s.data(), s.size());
    buf[s.size()] = 0;
  }
  return (int)s.size();
}
