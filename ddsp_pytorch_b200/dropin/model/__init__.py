"""Put this directory BEFORE the reference checkout on sys.path: ``model.ddsp.*`` then resolves to the B200
modules while ``model.autoencoder.*`` (untouched reference files) still resolves to the reference."""
from pkgutil import extend_path

__path__ = extend_path(__path__, __name__)
