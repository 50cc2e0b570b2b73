"""Import shim: the product package lives in the directory ``enarf-gan_b200/`` (the name the
build contract fixes).  A hyphen is not importable, so this module turns itself into that
package: ``import enarf_gan_b200`` / ``import enarf_gan_b200.rendering`` resolve into it."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "enarf-gan_b200")]
__file__ = _os.path.join(__path__[0], "__init__.py")
with open(__file__) as _f:
    exec(compile(_f.read(), __file__, "exec"))
