"""ctypes view of include/m2kdetect.h (struct layouts and constants)."""
import ctypes as C

NONE = 0xFFFFFFFF

OK, ERR_ARG, ERR_CUDA, ERR_IO, ERR_NOMEM, ERR_UNSUPPORTED, ERR_LIMIT, ERR_NOTDIR = 0, -1, -2, -3, -4, -5, -6, -7

T_ERR, T_REG, T_DIR, T_LNK, T_OTHER = 0, 1, 2, 3, 4
SH_BASH, SH_DASH = 0, 1

(F_PIPFILE, F_BUILD_XML, F_BUILD_GRADLE, F_POM_XML, F_PACKAGE_JSON, F_REQUIREMENTS, F_SETUP_PY,
 F_ENVIRONMENT_YML, F_GEMFILE, F_GO_MOD, R_GO, R_PHP, R_JAVA, N_WAR, N_IGNOREFILE, N_USER_DF,
 N_USER_S2I, N_YAML, C_MAIN_PY, C_DOCKERFILE) = range(20)

DF_NAMES = ["django", "golang", "java-war-jboss", "java-war-liberty", "java-war-tomcat", "javaant",
            "javagradle", "javamaven", "nodejs", "php", "python", "ruby"]
S2I_NAMES = ["golang", "java", "nodejs", "php", "python", "ruby"]
N_DF, N_S2I = len(DF_NAMES), len(S2I_NAMES)
DF_MASK, S2I_MASK = 0x00000FFF, 0x0003F000

FILE_HAS_MAIN, FILE_LONG_LINE, FILE_FROM_FIRST, FILE_DOCKERFILE, FILE_UNDECIDED = 1, 2, 4, 8, 16

(DIR_NOT_VISITED, DIR_PRECONTAINERIZED, DIR_IGNORED_PRUNE, DIR_IGNORED_DESCEND, DIR_NOMATCH_PRUNE,
 DIR_NOMATCH_DESCEND, DIR_SERVICE, DIR_WALK_ERROR) = range(8)
DIR_STATE_NAMES = ["not_visited", "precontainerized_prune", "ignored_prune", "ignored_descend",
                   "nomatch_prune", "nomatch_descend", "service", "walk_error"]

u8p, u32p, u64p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)


class Tree(C.Structure):
    _fields_ = [
        ("n_dirs", C.c_uint32), ("n_entries", C.c_uint32), ("n_files", C.c_uint32), ("max_depth", C.c_uint32),
        ("n_bytes", C.c_uint64), ("n_name_bytes", C.c_uint64), ("content_bytes", C.c_uint64),
        ("root_path", C.c_char_p),
        ("dir_parent", u32p), ("dir_end", u32p), ("dir_entry", u32p), ("dir_first_entry", u32p),
        ("dir_depth", u32p), ("dir_flags", u8p),
        ("entry_dir", u32p), ("entry_name_off", u32p), ("entry_type", u8p), ("entry_child", u32p),
        ("names", u8p),
        ("file_entry", u32p), ("file_off", u64p), ("file_len", u64p),
        ("bytes", u8p), ("owner", C.c_void_p),
    ]


class Result(C.Structure):
    _fields_ = [
        ("n_dirs", C.c_uint32), ("n_files", C.c_uint32), ("n_services", C.c_uint32),
        ("n_entries", C.c_uint32),
        ("dir_opts", u32p), ("dir_direct", u32p), ("dir_subtree", u32p), ("dir_state", u8p),
        ("dir_ignore", u8p), ("dir_main_file", u32p), ("file_flags", u8p), ("service_dirs", u32p), ("entry_feat", u32p),
        ("n_ignore_dirs", C.c_uint32), ("n_ignore_contents", C.c_uint32),
        ("ignore_dirs", C.POINTER(C.c_char_p)), ("ignore_contents", C.POINTER(C.c_char_p)),
        ("ms_h2d", C.c_float), ("ms_kernels", C.c_float), ("ms_scan_kernel", C.c_float),
        ("ms_d2h", C.c_float), ("ms_host_ignore", C.c_float), ("ms_total", C.c_float),
        ("n_kernel_launches", C.c_uint32), ("_pad", C.c_uint32),
        ("owner", C.c_void_p),
    ]


class DetectParams(C.Structure):
    _fields_ = [
        ("sh_flavor", C.c_uint32), ("n_pre", C.c_uint32),
        ("pre_paths", C.POINTER(C.c_char_p)),
        ("skip_content", C.c_uint32), ("_pad", C.c_uint32),
    ]


class SynthParams(C.Structure):
    _fields_ = [
        ("seed", C.c_uint64),
        ("n_files", C.c_uint32), ("n_dirs", C.c_uint32), ("max_depth", C.c_uint32), ("n_ignore_files", C.c_uint32),
        ("total_bytes", C.c_uint64),
        ("n_huge", C.c_uint32), ("_pad", C.c_uint32),
        ("huge_bytes", C.c_uint64),
        ("tiny_min", C.c_uint32), ("tiny_max", C.c_uint32),
        ("n_large", C.c_uint32), ("_pad2", C.c_uint32),
        ("large_bytes", C.c_uint64),
    ]
