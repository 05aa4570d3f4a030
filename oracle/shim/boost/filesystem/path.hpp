/* TEST INFRASTRUCTURE ONLY (oracle/_ref build): the one member the blocks use. */
#ifndef GB_ORACLE_SHIM_BOOST_FS_PATH_HPP
#define GB_ORACLE_SHIM_BOOST_FS_PATH_HPP
namespace boost { namespace filesystem { struct path { static constexpr char preferred_separator = '/'; }; } }
#endif
