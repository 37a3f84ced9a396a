from .Base import Base, MLPNode
from .SCFStack import SCFStack
from .EGCLStack import EGCLStack
from .PNAStack import PNAStack
from .PNAPlusStack import PNAPlusStack
from .PNAEqStack import PNAEqStack
from .PAINNStack import PAINNStack
from .create import create_model, create_model_config
