"""Build-time shim for installing the UNMODIFIED 2019 reference on Python 3.12.

The reference's `versioneer.py` (build tooling, not library code) calls
`configparser.SafeConfigParser().readfp(...)`, both removed in Python 3.12.  This file is put on
PYTHONPATH only while `baseline/install_ref.sh` runs pip; nothing under /root/reference is edited
and nothing of it is needed at run time."""
import configparser

if not hasattr(configparser, "SafeConfigParser"):
    class SafeConfigParser(configparser.ConfigParser):
        def readfp(self, fp, filename=None):
            return self.read_file(fp, filename)

    configparser.SafeConfigParser = SafeConfigParser
